// Version and error strings of the C ABI (include/vfdepth_b200.h).
#include "common.cuh"

extern "C" int vfd_version(void) { return VFD_VERSION; }

extern "C" const char* vfd_last_error(int code) {
    switch (code) {
        case VFD_OK: return "ok";
        case VFD_ERR_NULL: return "a required pointer is NULL";
        case VFD_ERR_DESC: return "descriptor field out of range";
        case VFD_ERR_ALIGN: return "pointer or leading dimension violates the 16-byte alignment contract";
        case VFD_ERR_WORKSPACE: return "workspace too small";
        case VFD_ERR_CUDA: return "CUDA reported an error at launch";
        case VFD_ERR_UNSUPPORTED: return "configuration not supported by this build";
        case VFD_ERR_STATIC_LIMIT:
            return "size exceeds the fixed capacity of this no-allocation entry point's static device scratch";
        default: return "unknown vfd_status";
    }
}
