// libRAFTStereoAlgorithm.so replacement (RAFTStereo/include/RAFTStereoAlgorithm.h:24-39): RunRAFTStereo always rectifies
// (RAFTStereo/src/RAFTStereoAlgorithm.cpp:57-72)
#define SA_SHIM_MODEL SA_MODEL_RAFTSTEREO
#define SA_SHIM_RUN_RECTIFY RunRAFTStereo
#include "stereo_algorithm_shim.h"
