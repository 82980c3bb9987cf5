// libHitNetAlgorithm.so replacement (HitNet/include/HitNetAlgorithm.h:24-39): RunHitNet always rectifies
// (HitNet/src/HitNetAlgorithm.cpp:40-50)
#define SA_SHIM_MODEL SA_MODEL_HITNET
#define SA_SHIM_RUN_RECTIFY RunHitNet
#include "stereo_algorithm_shim.h"
