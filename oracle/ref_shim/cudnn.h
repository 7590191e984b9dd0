// TEST INFRASTRUCTURE ONLY -- just enough of cudnn.h for LIB/include/cuda_lib/max_pool.hpp to
// parse.  The reference build under oracle/_ref does not link max_pool.cu; Pooling::execute is
// restated on the CPU in ref_api.cpp (integer max over disjoint windows, max_pool.cu:21-79).
#pragma once
typedef struct cudnnContext* cudnnHandle_t;
typedef struct cudnnTensorStruct* cudnnTensorDescriptor_t;
typedef struct cudnnPoolingStruct* cudnnPoolingDescriptor_t;
typedef enum { CUDNN_STATUS_SUCCESS = 0 } cudnnStatus_t;
typedef enum { CUDNN_DATA_INT8 = 3 } cudnnDataType_t;
inline const char* cudnnGetErrorString(cudnnStatus_t) { return "cudnn shim"; }
