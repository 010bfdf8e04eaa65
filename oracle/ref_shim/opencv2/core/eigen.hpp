// TEST INFRASTRUCTURE ONLY (oracle/): cv::eigen2cv for the CV_64F shim Mat.
#pragma once
#include <Eigen/Core>
#include <opencv2/opencv.hpp>
namespace cv
{
template <typename Derived>
void eigen2cv(const Eigen::MatrixBase<Derived>& src, Mat& dst)
{
	dst = Mat(static_cast<int>(src.rows()), static_cast<int>(src.cols()), CV_64F);
	for (int r = 0; r < src.rows(); ++r)
		for (int c = 0; c < src.cols(); ++c)
			dst.at<double>(r, c) = static_cast<double>(src(r, c));
}
}  // namespace cv
