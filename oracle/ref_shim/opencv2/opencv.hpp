// TEST INFRASTRUCTURE ONLY (oracle/): minimal stand-in for the slice of the
// OpenCV C++ API that the reference's feature_tracker/src/{patch,optimizer}.cpp
// and common/include/common/{geometry,data_types}.h touch.  OpenCV C++ is a
// system package of the reference (Dockerfile:24), not vendored, and is absent
// from this image, so the *unmodified* reference sources are compiled against
// this header instead.  Written from the documented OpenCV semantics, no
// OpenCV source is copied.  CV_64F matrices only.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <list>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_64F 6

namespace cv
{
// cvRound semantics: round half to even (SSE cvtsd2si under default MXCSR).
static inline int shimRound(double v)
{
	return static_cast<int>(std::nearbyint(v));
}

template <typename T>
struct shim_cast
{
	template <typename U>
	static T from(U v)
	{
		return static_cast<T>(v);
	}
};
template <>
struct shim_cast<int>
{
	static int from(double v) { return shimRound(v); }
	static int from(float v) { return shimRound(static_cast<double>(v)); }
	static int from(int v) { return v; }
};

template <typename T>
struct Point_
{
	T x, y;
	Point_() : x(0), y(0) {}
	// like cv::Point_(_Tp,_Tp): arguments convert by the language rules, so
	// Point2i(double,double) truncates.
	Point_(T x_, T y_) : x(x_), y(y_) {}
	// converting copy: saturate_cast (rounds for float->int).
	template <typename U>
	Point_(const Point_<U>& o)
		: x(shim_cast<T>::from(o.x)), y(shim_cast<T>::from(o.y))
	{
	}
};
template <typename T>
Point_<T> operator+(const Point_<T>& a, const Point_<T>& b)
{
	return Point_<T>(a.x + b.x, a.y + b.y);
}
template <typename T>
Point_<T> operator-(const Point_<T>& a, const Point_<T>& b)
{
	return Point_<T>(a.x - b.x, a.y - b.y);
}
template <typename T>
Point_<T> operator*(const Point_<T>& a, double s)
{
	return Point_<T>(shim_cast<T>::from(a.x * s), shim_cast<T>::from(a.y * s));
}
template <typename T>
Point_<T> operator*(double s, const Point_<T>& a)
{
	return a * s;
}
using Point2i = Point_<int>;
using Point = Point2i;
using Point2d = Point_<double>;
using Point2f = Point_<float>;

template <typename T>
struct Point3_
{
	T x, y, z;
	Point3_() : x(0), y(0), z(0) {}
	Point3_(T a, T b, T c) : x(a), y(b), z(c) {}
};
using Point3d = Point3_<double>;

template <typename T>
struct Size_
{
	T width, height;
	Size_() : width(0), height(0) {}
	Size_(T w, T h) : width(w), height(h) {}
};
using Size2i = Size_<int>;
using Size = Size2i;

template <typename T>
struct Rect_
{
	T x, y, width, height;
	Rect_() : x(0), y(0), width(0), height(0) {}
	Rect_(T x_, T y_, T w_, T h_) : x(x_), y(y_), width(w_), height(h_) {}
	template <typename U>
	Rect_(const Rect_<U>& o)
		: x(shim_cast<T>::from(o.x)),
		  y(shim_cast<T>::from(o.y)),
		  width(shim_cast<T>::from(o.width)),
		  height(shim_cast<T>::from(o.height))
	{
	}
	Point_<T> tl() const { return Point_<T>(x, y); }
	Point_<T> br() const { return Point_<T>(x + width, y + height); }
	bool contains(const Point_<T>& p) const
	{
		return x <= p.x && p.x < x + width && y <= p.y && p.y < y + height;
	}
};
using Rect2d = Rect_<double>;
using Rect = Rect_<int>;

enum
{
	NORM_L1 = 2,
	NORM_L2 = 4
};
enum
{
	INTER_NEAREST = 0,
	INTER_LINEAR = 1,
	WARP_INVERSE_MAP = 16
};

class Mat
{
   public:
	int rows = 0, cols = 0;
	unsigned char* data = nullptr;

	Mat() {}
	Mat(int r, int c, int /*type*/) { create(r, c); }
	// header over caller-owned pixels, cv::Mat(rows, cols, type, data): no copy, as in
	// OpenCV (how a CV_8U frame reaches FeatureDetector::newImage in the drop-in tests)
	Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data(static_cast<unsigned char*>(ext)), type_(type), step_(c) {}

	static Mat zeros(int r, int c, int type)
	{
		Mat m(r, c, type);
		return m;
	}
	static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }

	void create(int r, int c)
	{
		rows = r;
		cols = c;
		step_ = c;
		buf_ = std::make_shared<std::vector<double>>(
			static_cast<size_t>(r) * c, 0.0);
		base_ = buf_->data();
		data = reinterpret_cast<unsigned char*>(base_);
	}
	int type() const { return type_; }
	int channels() const { return 1; }
	bool isContinuous() const { return step_ == cols; }
	bool empty() const { return rows == 0 || cols == 0; }

	template <typename T>
	T& at(int y, int x)
	{
		return base_[static_cast<size_t>(y) * step_ + x];
	}
	template <typename T>
	const T& at(int y, int x) const
	{
		return base_[static_cast<size_t>(y) * step_ + x];
	}

	// ROI view sharing the buffer (argument converts Rect2d -> Rect by
	// rounding, as in OpenCV).
	Mat operator()(const Rect& r) const
	{
		Mat v;
		v.rows = r.height;
		v.cols = r.width;
		v.step_ = step_;
		v.buf_ = buf_;
		v.base_ = base_ + static_cast<size_t>(r.y) * step_ + r.x;
		v.data = reinterpret_cast<unsigned char*>(v.base_);
		return v;
	}
	Mat clone() const
	{
		Mat m(rows, cols, CV_64F);
		for (int y = 0; y < rows; ++y)
			for (int x = 0; x < cols; ++x) m.at<double>(y, x) = at<double>(y, x);
		return m;
	}

	// 8-bit image that owns its pixels (what cv::imread returns)
	static Mat owned8(int r, int c, const unsigned char* px)
	{
		Mat m;
		m.rows = r;
		m.cols = c;
		m.type_ = CV_8U;
		m.step_ = c;
		m.bytes_ = std::make_shared<std::vector<unsigned char>>(px, px + static_cast<size_t>(r) * c);
		m.data = m.bytes_->data();
		return m;
	}

   private:
	int type_ = CV_64F;
	int step_ = 0;
	std::shared_ptr<std::vector<double>> buf_;
	std::shared_ptr<std::vector<unsigned char>> bytes_;
	double* base_ = nullptr;
};

// cv::imread(path, flags) for the one format the replay test writes: binary PGM ("P5", maxval
// 255), whatever the file is called (the reference's images.txt names .png files; OpenCV picks
// the decoder by content as well).  An unreadable file gives an empty Mat, as in OpenCV.
static inline Mat imread(const std::string& path, int /*flags*/ = 0)
{
	FILE* f = std::fopen(path.c_str(), "rb");
	if (!f) return Mat();
	int w = 0, h = 0, maxv = 0;
	char magic[3] = {0, 0, 0};
	Mat out;
	if (std::fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) == 4 && magic[0] == 'P' && magic[1] == '5' && maxv == 255 &&
		w > 0 && h > 0 && std::fgetc(f) != EOF)
	{
		std::vector<unsigned char> px(static_cast<size_t>(w) * h);
		if (std::fread(px.data(), 1, px.size(), f) == px.size()) out = Mat::owned8(h, w, px.data());
	}
	std::fclose(f);
	return out;
}

// dst = a*alpha + b*beta  (the evaluation OpenCV's lazy MatExpr performs for
// scaled sums; b may be empty).
static inline Mat shimAxpby(const Mat& a, double alpha, const Mat* b,
							double beta)
{
	Mat m(a.rows, a.cols, CV_64F);
	for (int y = 0; y < a.rows; ++y)
		for (int x = 0; x < a.cols; ++x)
			m.at<double>(y, x) =
				b ? a.at<double>(y, x) * alpha + b->at<double>(y, x) * beta
				  : a.at<double>(y, x) * alpha;
	return m;
}
static inline Mat operator*(const Mat& a, double s)
{
	return shimAxpby(a, s, nullptr, 0);
}
static inline Mat operator*(double s, const Mat& a)
{
	return shimAxpby(a, s, nullptr, 0);
}
// OpenCV evaluates Mat / scalar as a scale by the reciprocal.
static inline Mat operator/(const Mat& a, double s)
{
	return shimAxpby(a, 1.0 / s, nullptr, 0);
}
static inline Mat operator-(const Mat& a)
{
	return shimAxpby(a, -1.0, nullptr, 0);
}
static inline Mat operator+(const Mat& a, const Mat& b)
{
	return shimAxpby(a, 1.0, &b, 1.0);
}
static inline Mat operator-(const Mat& a, const Mat& b)
{
	return shimAxpby(a, 1.0, &b, -1.0);
}
static inline Mat operator+(const Mat& a, double s)
{
	Mat m = a.clone();
	for (int y = 0; y < m.rows; ++y)
		for (int x = 0; x < m.cols; ++x) m.at<double>(y, x) += s;
	return m;
}

static inline double norm(const Mat& m, int type = NORM_L2)
{
	double acc = 0;
	for (int y = 0; y < m.rows; ++y)
		for (int x = 0; x < m.cols; ++x)
		{
			const double v = m.at<double>(y, x);
			acc += (type == NORM_L1) ? std::fabs(v) : v * v;
		}
	return type == NORM_L1 ? acc : std::sqrt(acc);
}
template <typename T>
double norm(const Point_<T>& p)
{
	return std::sqrt(static_cast<double>(p.x) * p.x +
					 static_cast<double>(p.y) * p.y);
}

// Classic (pre-4.x and 4.x non-"algorithm hint") warpAffine, nearest
// neighbour, constant-0 border, M used directly as dst->src map when
// WARP_INVERSE_MAP is set: 10-bit fixed-point coordinates with
// round-half-even of each addend (verified there
// against cv2 4.13 float32).  Without WARP_INVERSE_MAP the matrix is inverted
// first, as OpenCV does.
static inline void warpAffine(const Mat& src, Mat& dst, const Mat& M, Size sz,
							  int flags)
{
	double m[6] = {M.at<double>(0, 0), M.at<double>(0, 1), M.at<double>(0, 2),
				   M.at<double>(1, 0), M.at<double>(1, 1), M.at<double>(1, 2)};
	if (!(flags & WARP_INVERSE_MAP))
	{
		double D = m[0] * m[4] - m[1] * m[3];
		D = D != 0 ? 1. / D : 0;
		const double A11 = m[4] * D, A22 = m[0] * D;
		m[0] = A11;
		m[1] *= -D;
		m[3] *= -D;
		m[4] = A22;
		const double b1 = -m[0] * m[2] - m[1] * m[5];
		const double b2 = -m[3] * m[2] - m[4] * m[5];
		m[2] = b1;
		m[5] = b2;
	}
	Mat out(sz.height, sz.width, CV_64F);
	for (int y = 0; y < sz.height; ++y)
		for (int x = 0; x < sz.width; ++x)
		{
			const int X0 = shimRound(m[0] * x * 1024) +
						   shimRound((m[1] * y + m[2]) * 1024) + 512;
			const int Y0 = shimRound(m[3] * x * 1024) +
						   shimRound((m[4] * y + m[5]) * 1024) + 512;
			const int xi = X0 >> 10, yi = Y0 >> 10;
			out.at<double>(y, x) =
				(xi >= 0 && xi < src.cols && yi >= 0 && yi < src.rows)
					? src.at<double>(yi, xi)
					: 0.0;
		}
	dst = out;
}
}  // namespace cv
