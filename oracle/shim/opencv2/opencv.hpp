// Empty stand-in for <opencv2/opencv.hpp>  --  TEST INFRASTRUCTURE, NOT PRODUCT.
// The reference's .cu files include the OpenCV umbrella header without using anything from it
// (e.g. FastACVNet_plus/src/FastACVNet_plus_preprocess.cu:2).  OpenCV's C++ headers are not in
// this image, so oracle/build_ref.py puts this directory on the include path to compile those
// files UNMODIFIED, from where they lie, into oracle/_ref/.
#pragma once
