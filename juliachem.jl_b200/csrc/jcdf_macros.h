// Status-returning error macros shared by the host translation units of libjcdf (api.cu, scf.cu, integrals.cu).
#pragma once
#include <string>
#include <cuda_runtime.h>

#include "jcdf_internal.h"
#include "nccl_dyn.h"

#define JC_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess) {                                                                       \
            set_error(std::string(#call) + ": " + cudaGetErrorString(e__) + " (" + __FILE__ + ":" +     \
                      std::to_string(__LINE__) + ")");                                                  \
            return (e__ == cudaErrorMemoryAllocation) ? JCDF_ERR_NOMEM : JCDF_ERR_CUDA;                 \
        }                                                                                               \
    } while (0)
#define JC_CUBLAS(call)                                                                                 \
    do {                                                                                                \
        cublasStatus_t s__ = (call);                                                                    \
        if (s__ != CUBLAS_STATUS_SUCCESS) {                                                             \
            set_error(std::string(#call) + ": cuBLAS status " + std::to_string((int)s__));              \
            return JCDF_ERR_CUDA;                                                                       \
        }                                                                                               \
    } while (0)
#define JC_CUSOLVER(call)                                                                               \
    do {                                                                                                \
        cusolverStatus_t s__ = (call);                                                                  \
        if (s__ != CUSOLVER_STATUS_SUCCESS) {                                                           \
            set_error(std::string(#call) + ": cuSOLVER status " + std::to_string((int)s__));            \
            return JCDF_ERR_CUDA;                                                                       \
        }                                                                                               \
    } while (0)
#define JC_NCCL(call)                                                                                   \
    do {                                                                                                \
        int r__ = (call);                                                                               \
        if (r__ != 0) {                                                                                 \
            set_error(std::string(#call) + ": " + nccl().GetErrorString(r__));                          \
            return JCDF_ERR_NCCL;                                                                       \
        }                                                                                               \
    } while (0)
#define JC_REQUIRE(cond, msg)                    \
    do {                                         \
        if (!(cond)) {                           \
            set_error(msg);                      \
            return JCDF_ERR_INVALID;             \
        }                                        \
    } while (0)

