// peer.cu -- the small exchanges of the several-rank summary as plain stores into peer memory.
//
// An all-gather of a few hundred KiB through NCCL costs 20-40 us of fixed latency per call, four
// times per aggregate step.  With every rank's receive buffer mapped into every process (symmetric
// memory: torch.distributed._symmetric_memory, CUDA VMM handles exchanged once), a rank "sends" by
// storing its block into slot `rank` of every peer's buffer over NVLink, 16 bytes per thread, and
// the ranks then meet at a device-side barrier; nothing else moves.  Only ordinary global stores
// are used (no TMA, no multicast).
#include "pmx_common.cuh"

namespace {

__global__ void __launch_bounds__(256)
peer_put_kernel(const uint4* __restrict__ src, const unsigned long long* __restrict__ peer_base, int world, int rank,
                unsigned long long dst_offset, unsigned long long n16) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n16) return;
  const uint4 v = src[i];
  for (int p = 0; p < world; ++p) {
    uint4* dst = reinterpret_cast<uint4*>(peer_base[p] + dst_offset) + (unsigned long long)rank * n16;
    dst[i] = v;
  }
}

}  // namespace

extern "C" int pmx_peer_put(const void* d_src, const uint64_t* d_peer_base, int world, int rank, uint64_t dst_offset_bytes,
                            uint64_t nbytes, pmx_stream stream) {
  if (world < 1 || rank < 0 || rank >= world) return pmx_fail(PMX_E_INVALID, "pmx_peer_put: bad rank / world");
  if (nbytes == 0) return PMX_OK;
  if (!d_src || !d_peer_base) return pmx_fail(PMX_E_INVALID, "pmx_peer_put: NULL buffer");
  if ((nbytes & 15) || (dst_offset_bytes & 15) || (reinterpret_cast<uintptr_t>(d_src) & 15))
    return pmx_fail(PMX_E_INVALID, "pmx_peer_put: sizes, offsets and the source must be multiples of 16 bytes");
  const unsigned long long n16 = nbytes / 16;
  peer_put_kernel<<<(unsigned)((n16 + 255) / 256), 256, 0, pmx_st(stream)>>>((const uint4*)d_src, (const unsigned long long*)d_peer_base,
                                                                            world, rank, dst_offset_bytes, n16);
  PMX_LAUNCH_CHECK();
  return PMX_OK;
}
