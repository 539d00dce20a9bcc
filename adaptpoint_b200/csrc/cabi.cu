// Library identification and status strings for libapb200.so.
#include "apb_common.cuh"

extern "C" const char *apb_version(void) { return "apb200 0.1.0 sm_100a"; }

extern "C" const char *apb_error_string(int status) {
    switch (status) {
        case APB_OK: return "ok";
        case APB_E_INVALID: return "apb200: invalid argument";
        case APB_E_UNSUPPORTED: return "apb200: shape not supported by this kernel family";
        case APB_E_SCRATCH_LIMIT: return "apb200: problem exceeds the library's built-in scratch array (use the entry point that takes a workspace)";
        default: return cudaGetErrorString((cudaError_t)status);
    }
}
