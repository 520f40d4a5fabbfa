// window_fast.cu -- the fused kernel on raw samples (align + rolling + rate), all window phases.
#include "window_fast_impl.cuh"

namespace pmxw {
template int launch_window_fast<true>(const pmx_ctx*, WinParams, cudaStream_t);
}
