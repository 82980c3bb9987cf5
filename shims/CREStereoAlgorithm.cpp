// libCREStereoAlgorithm.so replacement (CREStereo/include/CREStereoAlgorithm.h:24-44)
#define SA_SHIM_MODEL SA_MODEL_CRESTEREO
#define SA_SHIM_RUN RunCREStereo
#define SA_SHIM_RUN_RECTIFY RunCREStereo_RectifyImage
#include "stereo_algorithm_shim.h"
