// libFastACVNet_plus_Algorithm.so replacement (FastACVNet_plus/include/FastACVNet_plus_Algorithm.h:23-43)
#define SA_SHIM_MODEL SA_MODEL_FASTACVNET_PLUS
#define SA_SHIM_RUN RunFastACVNet_plus
#define SA_SHIM_RUN_RECTIFY RunFastACVNet_plus_RectifyImage
#include "stereo_algorithm_shim.h"
