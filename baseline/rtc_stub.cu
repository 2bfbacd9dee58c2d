// baseline/rtc_stub.cu - tiny-cuda-nn's no-RTC branch (src/rtc_kernel.cu:63-78) omits CudaRtcKernel::set, which the
// headers still reference from the JIT paths (never taken here: the renderer leaves jit_fusion off, object.h:761).
// The reference's real build links nvrtc + a CMake-generated resource library instead; this one-symbol stub lets the
// direct nvcc recipe of baseline/Makefile link without either.
#include <tiny-cuda-nn/rtc_kernel.h>
namespace tcnn {
void CudaRtcKernel::set(CUfunction_attribute, int) {}
}
