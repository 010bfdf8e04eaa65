// TEST / BENCH INFRASTRUCTURE ONLY — CPU restatement of the reference's event
// text reader, used as the parser's CPU baseline and large-input checker.
//
// Follows tools::Davis240cReader::getEventSample line by line
// (/root/reference/tools/dataset_reader/src/davis240c_reader.cpp:60-92) and the
// std::getline loop of DatasetReader::readFile: std::string::find / substr,
// std::stod, duration_cast<microseconds>(duration<double>), std::stoi, the
// 0/1 sign check.  The reference file itself needs OpenCV (cv::imread) and
// spdlog and is not compilable here.
#include <chrono>
#include <cstdint>
#include <sstream>
#include <stdexcept>
#include <string>

#include "../include/ebt_types.h"

extern "C" int64_t ingest_parse_text(const char* text, size_t len, ebt_event* out, int64_t cap)
{
	using timestamp_t = std::chrono::microseconds;
	std::istringstream in(std::string(text, len));
	std::string line;
	int64_t n = 0;
	try
	{
		while (std::getline(in, line))
		{
			if (n >= cap) return -2;
			size_t pos = line.find(' ');
			const auto duration = std::chrono::duration<double>(std::stod(line.substr(0, pos)));
			const timestamp_t timestamp = std::chrono::duration_cast<timestamp_t>(duration);
			line = line.substr(pos + 1);
			pos = line.find(' ');
			const int x = std::stoi(line.substr(0, pos));
			line = line.substr(pos + 1);
			pos = line.find(' ');
			const int y = std::stoi(line.substr(0, pos));
			line = line.substr(pos + 1);
			const int sign = std::stoi(line.substr(0, pos));
			if (sign != 0 && sign != 1) throw std::runtime_error("Sign is not equal to 0/1");
			ebt_event e{};
			e.t_us = timestamp.count();
			e.x = static_cast<uint16_t>(x);
			e.y = static_cast<uint16_t>(y);
			e.polarity = sign ? 1 : -1;
			out[n++] = e;
		}
	}
	catch (const std::exception&)
	{
		return -1;
	}
	return n;
}
