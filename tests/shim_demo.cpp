// A cut-down FastACVNet_plus/test/main.cpp: same calls, same order, raw files instead of imread/imwrite.
//   shim_demo <calibration.yml> <left.bgr> <right.bgr> <W> <H> <out_prefix>
#include <cstdio>
#include <fstream>
#include <vector>
#include <opencv2/opencv.hpp>

extern "C" void* Initialize(char* model_path, int gpu_id, char* calibration_path);
extern "C" int RunFastACVNet_plus_RectifyImage(void* p, cv::Mat& left_image, cv::Mat& right_image, float* pointcloud, cv::Mat& disparity);
extern "C" const char* Version(void* p);
extern "C" int Release(void* p);

static bool read_file(const char* path, unsigned char* dst, size_t n)
{
    std::ifstream f(path, std::ios::binary);
    f.read(reinterpret_cast<char*>(dst), (std::streamsize)n);
    return (size_t)f.gcount() == n;
}
static void write_file(const std::string& path, const void* src, size_t n)
{
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<const char*>(src), (std::streamsize)n);
}

int main(int argc, char** argv)
{
    if (argc < 7) return 2;
    const int W = std::atoi(argv[4]), H = std::atoi(argv[5]);
    const std::string out = argv[6];
    cv::Mat imageL(H, W, CV_8UC3), imageR(H, W, CV_8UC3);
    if (!read_file(argv[2], imageL.data, (size_t)W * H * 3) || !read_file(argv[3], imageR.data, (size_t)W * H * 3)) return 3;
    char model[] = "fast_acvnet_plus_generalization_opset16_480x640.onnx";
    void* fastacvnet = Initialize(model, 0, argv[1]);
    std::printf("%s\n", Version(fastacvnet));
    std::vector<float> pointcloud((size_t)W * H * 6);
    cv::Mat disparity;
    cv::Mat imageL1, imageR1;
    for (int i = 0; i < 5; i++) {
        imageL1 = imageL.clone();
        imageR1 = imageR.clone();
        if (RunFastACVNet_plus_RectifyImage(fastacvnet, imageL1, imageR1, pointcloud.data(), disparity) != 0) return 4;
    }
    write_file(out + "_rect_left.bgr", imageL1.data, (size_t)W * H * 3);
    write_file(out + "_rect_right.bgr", imageR1.data, (size_t)W * H * 3);
    write_file(out + "_disparity.f32", disparity.data, (size_t)W * H * 4);
    write_file(out + "_pointcloud.f32", pointcloud.data(), pointcloud.size() * 4);
    if (disparity.rows != H || disparity.cols != W || disparity.type() != CV_32FC1) return 5;
    return Release(fastacvnet);
}
