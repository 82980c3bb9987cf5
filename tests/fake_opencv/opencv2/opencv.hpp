// Minimal stand-in for the parts of cv::Mat the shims touch, so that shims/*.cpp can be compiled and run in an image
// without OpenCV's C++ headers.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <memory>

#define CV_8UC3 16
#define CV_32FC1 5

namespace cv {
class Mat {
public:
    int rows = 0, cols = 0;
    unsigned char* data = nullptr;
    size_t step[2] = {0, 0};
    Mat() {}
    Mat(int r, int c, int t) { create(r, c, t); }
    void create(int r, int c, int t)
    {
        type_ = t;
        rows = r; cols = c;
        const size_t es = t == CV_8UC3 ? 3 : 4;
        step[0] = (size_t)c * es; step[1] = es;
        buf_.reset(new unsigned char[(size_t)r * step[0]], std::default_delete<unsigned char[]>());
        data = buf_.get();
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    Mat clone() const { Mat m; if (!empty()) { m.create(rows, cols, type_); std::memcpy(m.data, data, (size_t)rows * step[0]); } return m; }
private:
    int type_ = 0;
    std::shared_ptr<unsigned char> buf_;
};
}  // namespace cv
