// pgen_attention_static: the attention pass of pgen_attention for callers that must not allocate.
//
// Who cannot allocate: an operator thread embedded in a host process whose device memory is wholly
// owned by somebody else's allocator (a framework pool sized to the card, a memory budget enforced by the
// task manager), and service loops that may not take the stalls cudaMalloc / cudaFree bring (both
// serialise against the whole device and take the driver's allocation lock).  For them the energies
// scratch lives in a fixed file-scope device array instead of a per-call cudaMalloc.
//
// The array's name below is only a host-side stand-in for the device symbol: handing it to a kernel as
// if it were an address would fault (or, where the GPU can reach host memory, quietly use the wrong
// bytes).  The real device address is looked up with cudaGetSymbolAddress, once per device, with that
// device current, and that pointer is what the kernels receive.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "handle_access.h"
#include "kernels.h"

namespace {

constexpr size_t kScratchBytes = PGEN_ATTENTION_STATIC_SCRATCH_BYTES;   // 1 MiB, see include/pgen.h
constexpr int kMaxDevices = 64;

// Held by every GPU that loads the library, until the process exits.
static __device__ __align__(256) unsigned char g_attention_scratch[kScratchBytes];

std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {};   // device address of g_attention_scratch, per device ordinal

// Resolves the array's address on `device` (which must be current).  Runs the lookup the first time a
// device is seen; later calls return the cached pointer.
cudaError_t scratch_address(int device, void** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_addr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_attention_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_addr[device] = p;
  }
  *out = g_scratch_addr[device];
  return cudaSuccess;
}

__global__ void mark_articles_active_kernel(int* flags, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = 1;
}

size_t round_up_256(size_t x) { return (x + 255) / 256 * 256; }

}  // namespace

extern "C" int pgen_attention_static(pgen_t* h, int n_articles, const float* enc_feat, const float* enc_states,
                                     const int32_t* lens, const float* dfeat, const float* cov, float* attn,
                                     float* ctx) {
  if (!h) return PGEN_ERR_ARG;
  if (n_articles < 0 || (n_articles && (!enc_feat || !enc_states || !lens || !dfeat || !attn || !ctx)))
    return pgen_internal_fail(h, PGEN_ERR_ARG, "pgen_attention_static: NULL device pointer");
  if (n_articles && pgen_internal_coverage(h) && !cov)
    return pgen_internal_fail(h, PGEN_ERR_ARG, "pgen_attention_static: coverage=1 needs cov");
  if (n_articles == 0) return PGEN_OK;
  pgn::Engine* eng = pgen_internal_engine(h);
  pgn::Dims d = eng->dims();
  // scratch layout: [ one "article is live" flag per article | energies: one float per row and position ]
  const size_t flag_bytes = round_up_256((size_t)n_articles * sizeof(int));
  const size_t need = flag_bytes + (size_t)n_articles * d.K * d.T * sizeof(float);
  if (need > kScratchBytes)
    return pgen_internal_fail(h, PGEN_ERR_ARG,
        "pgen_attention_static: " + std::to_string(n_articles) + " articles x beam " + std::to_string(d.K) +
        " x " + std::to_string(d.T) + " positions need " + std::to_string(need) + " bytes of scratch, the static array holds " +
        std::to_string(kScratchBytes) + "; use pgen_attention or fewer articles per call");
  const int device = eng->device();
  if (device < 0 || device >= kMaxDevices)
    return pgen_internal_fail(h, PGEN_ERR_ARG, "pgen_attention_static: device ordinal beyond the per-device table");
  cudaError_t ce = cudaSetDevice(device);
  void* base = nullptr;
  if (ce == cudaSuccess) ce = scratch_address(device, &base);
  if (ce != cudaSuccess)
    return pgen_internal_fail(h, PGEN_ERR_CUDA, std::string("pgen_attention_static set-up: ") + cudaGetErrorString(ce));
  int* flags = reinterpret_cast<int*>(base);
  float* energies = reinterpret_cast<float*>(static_cast<unsigned char*>(base) + flag_bytes);
  cudaStream_t st = eng->stream();
  // The array keeps whatever the previous call left in it.  The flags are rewritten here; the energies
  // of every valid (row, position) are written by the first kernel before the second one reads them,
  // so nothing else has to start from zero.
  mark_articles_active_kernel<<<(n_articles + 127) / 128, 128, 0, st>>>(flags, n_articles);
  d.SW = d.A;                                     // ctx rows are dense [rows, 2H]
  const pgn::DeviceWeights& w = eng->weights();
  eng->stats.kernel_launches += 1;
  eng->stats.kernel_launches += pgn::launch_attn_energy(d, n_articles, d.K, lens, flags, enc_feat, dfeat, cov, w.v, w.w_c, energies, st);
  eng->stats.kernel_launches += pgn::launch_attn_context(d, n_articles, d.K, lens, flags, enc_states, energies, attn, ctx, st);
  ce = cudaStreamSynchronize(st);
  if (ce == cudaSuccess) ce = cudaGetLastError();
  if (ce != cudaSuccess)
    return pgen_internal_fail(h, PGEN_ERR_CUDA, std::string("pgen_attention_static: ") + cudaGetErrorString(ce));
  return PGEN_OK;
}
