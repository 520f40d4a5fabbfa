// window_fast_dense.cu -- the same kernel on an already aligned [S,T] matrix (pmx_rolling / pmx_rate).
#include "window_fast_impl.cuh"

namespace pmxw {
template int launch_window_fast<false>(const pmx_ctx*, WinParams, cudaStream_t);
}
