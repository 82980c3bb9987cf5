// Source-level drop-in for the reference's per-model shared libraries.
//
// The reference builds, per model, a library that exports  Initialize / Run<Model>[_RectifyImage] / Version / Release
// with cv::Mat& parameters (e.g. FastACVNet_plus/include/FastACVNet_plus_Algorithm.h:23-43).  Compile ONE of the four
// .cpp files next to this header - against the application's own OpenCV - into a library of the same name and link it
// with libstereo_geom.so: the application keeps including the reference's header and calling the reference's names.
// Everything here is a thin adapter from cv::Mat to the C ABI in include/stereo_algorithms_compat.h.
//
// A translation unit defines, before including this file:
//   SA_SHIM_MODEL        one of SA_MODEL_RAFTSTEREO / HITNET / CRESTEREO / FASTACVNET_PLUS
//   SA_SHIM_RUN          name of the entry point that does NOT rectify (CREStereo, FastACVNet_plus), or nothing
//   SA_SHIM_RUN_RECTIFY  name of the entry point that rectifies first
#pragma once
#include <opencv2/opencv.hpp>

#include "../include/stereo_algorithms_compat.h"

namespace sa_shim {

inline sa_image as_image(cv::Mat& m)
{
    sa_image im;
    im.data = m.empty() ? nullptr : m.data;
    im.rows = m.rows;
    im.cols = m.cols;
    im.step = m.empty() ? 0 : (size_t)m.step[0];
    return im;
}

inline int run(void* p, int rectify, cv::Mat& left, cv::Mat& right, float* pointcloud, cv::Mat& disparity)
{
    if (!left.empty() && left.type() != CV_8UC3) return -1;
    if (!right.empty() && right.type() != CV_8UC3) return -1;
    sa_image l = as_image(left), r = as_image(right);
    cv::Mat d;
    if (!left.empty()) d.create(left.rows, left.cols, CV_32FC1);       // continuous
    const int rc = sa_ref_Run(p, rectify, &l, &r, pointcloud, d.empty() ? nullptr : reinterpret_cast<float*>(d.data));
    if (rc == 0) disparity = d;   // DisparityMap = DisparityImage.clone() (FastACVNet_plus/src/TRTFastACVNet_plus.cpp:167-168)
    return rc;
}

}  // namespace sa_shim

extern "C" {

void* Initialize(char* model_path, int gpu_id, char* calibration_path) { return sa_ref_Initialize(SA_SHIM_MODEL, model_path, gpu_id, calibration_path); }

#ifdef SA_SHIM_RUN
int SA_SHIM_RUN(void* p, cv::Mat& left_image, cv::Mat& right_image, float* pointcloud, cv::Mat& disparity)
{
    return sa_shim::run(p, 0, left_image, right_image, pointcloud, disparity);
}
#endif

int SA_SHIM_RUN_RECTIFY(void* p, cv::Mat& left_image, cv::Mat& right_image, float* pointcloud, cv::Mat& disparity)
{
    return sa_shim::run(p, 1, left_image, right_image, pointcloud, disparity);
}

const char* Version(void* p) { return sa_ref_Version(p); }

int Release(void* p) { return sa_ref_Release(p); }

}  // extern "C"
