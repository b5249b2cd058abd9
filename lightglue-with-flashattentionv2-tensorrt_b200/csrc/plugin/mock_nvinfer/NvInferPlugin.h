// MOCK companion of <NvInferPlugin.h>; everything the plugin TU needs is in the mock NvInfer.h.
#ifndef LG_MOCK_NVINFER_PLUGIN_H_
#define LG_MOCK_NVINFER_PLUGIN_H_
#include "NvInfer.h"
#endif
