/* TEST DOUBLE of the CUDA runtime, for tests/test_host_logic_fake_cuda.py only.
 *
 * The library's host logic (csrc/capi.cu, csrc/engine.cu: row batching, slots, flush deadline, in-order results, what
 * happens to rows when a CUDA call fails) can otherwise only run on a GPU box, and its error paths not even there.
 * Here the library's own objects (the .o files under csrc/build, unmodified) are linked against this file instead of libcudart:
 * "device" memory is zeroed host memory, copies are memcpy, kernel launches and graphs do nothing -- so every decoded
 * summary is empty -- and any CUDA call can be told to fail (fake_cuda_fail).  Nothing computes; nothing ships; the
 * product never links this (csrc/build.sh links the real runtime, statically). */
#include <cuda_runtime_api.h>
#include <stdlib.h>
#include <string.h>

#define API __attribute__((visibility("default")))

/* like the real runtime: current device, last error and the pending call configuration are per THREAD; the counters
 * and the logs are shared and updated atomically / under a lock, so that tools/host_tsan.sh reports the library's races,
 * not this file's */
#include <pthread.h>
static int g_devices = 2;
static __thread int g_device = 0;
static long g_launches = 0, g_graph_launches = 0, g_live_allocs = 0;
static __thread cudaError_t g_last = cudaSuccess;
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
#define INC(x) __atomic_add_fetch(&(x), 1, __ATOMIC_RELAXED)
#define DEC(x) __atomic_sub_fetch(&(x), 1, __ATOMIC_RELAXED)
static char g_fail_name[64];
static int g_fail_after = -1;                     /* fail the call after this many more successful calls of that name */

API void fake_cuda_set_devices(int n) { g_devices = n; }
API void fake_cuda_fail(const char* name, int after) { strncpy(g_fail_name, name ? name : "", 63); g_fail_after = name ? after : -1; }
API long fake_cuda_launches(void) { return g_launches; }
API long fake_cuda_graph_launches(void) { return g_graph_launches; }
API long fake_cuda_live_allocs(void) { return g_live_allocs; }

static cudaError_t gate(const char* name) {
  if (__atomic_load_n(&g_fail_after, __ATOMIC_RELAXED) < 0) return cudaSuccess;
  cudaError_t r = cudaSuccess;
  pthread_mutex_lock(&g_mu);
  if (g_fail_after >= 0 && strcmp(name, g_fail_name) == 0) {
    if (g_fail_after-- == 0) { g_last = cudaErrorUnknown; r = cudaErrorUnknown; }
  }
  pthread_mutex_unlock(&g_mu);
  return r;
}
#define GATE(name) do { cudaError_t e__ = gate(name); if (e__ != cudaSuccess) return e__; } while (0)

API cudaError_t cudaGetDeviceCount(int* n) { *n = g_devices; return g_devices > 0 ? cudaSuccess : cudaErrorNoDevice; }
API cudaError_t cudaSetDevice(int d) { GATE("cudaSetDevice"); if (d < 0 || d >= g_devices) return cudaErrorInvalidDevice; g_device = d; return cudaSuccess; }
API cudaError_t cudaGetDevice(int* d) { *d = g_device; return cudaSuccess; }
API cudaError_t cudaDeviceSynchronize(void) { GATE("cudaDeviceSynchronize"); return cudaSuccess; }
API cudaError_t cudaDeviceGetAttribute(int* v, enum cudaDeviceAttr a, int dev) {
  *v = a == cudaDevAttrMultiProcessorCount ? 148 : a == cudaDevAttrMaxSharedMemoryPerBlockOptin ? 232448 : 0;
  return cudaSuccess;
}
API cudaError_t cudaGetLastError(void) { cudaError_t e = g_last; g_last = cudaSuccess; return e; }
API const char* cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "fake CUDA failure"; }

API cudaError_t cudaMalloc(void** p, size_t n) { GATE("cudaMalloc"); *p = calloc(n ? n : 1, 1); INC(g_live_allocs); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
API cudaError_t cudaFree(void* p) { if (p) { free(p); DEC(g_live_allocs); } return cudaSuccess; }
API cudaError_t cudaMallocHost(void** p, size_t n) { GATE("cudaMallocHost"); *p = calloc(n ? n : 1, 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
API cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
static int g_skip_async_memset = 0;               /* for timing the HOST cost of a batch: device-side fills cost the host nothing */
API void fake_cuda_skip_async_memset(int on) { g_skip_async_memset = on; }
API cudaError_t cudaMemset(void* p, int v, size_t n) { memset(p, v, n); return cudaSuccess; }
API cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t s) { if (!g_skip_async_memset) memset(p, v, n); return cudaSuccess; }
API cudaError_t cudaMemcpy(void* d, const void* s, size_t n, enum cudaMemcpyKind k) { GATE("cudaMemcpy"); memmove(d, s, n); return cudaSuccess; }
/* the last host-to-device copies, so that a test can look at what the library uploaded (the "device" is host memory) */
static struct { void* dst; size_t n; } g_h2d[16];
static long g_h2d_count = 0;
API long fake_cuda_h2d_count(void) { return g_h2d_count; }
API void* fake_cuda_h2d_dst(long back) { return g_h2d[(g_h2d_count - 1 - back) & 15].dst; }
API size_t fake_cuda_h2d_bytes(long back) { return g_h2d[(g_h2d_count - 1 - back) & 15].n; }
API cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, enum cudaMemcpyKind k, cudaStream_t st) {
  GATE("cudaMemcpyAsync");
  memmove(d, s, n);
  if (k == cudaMemcpyHostToDevice) {
    pthread_mutex_lock(&g_mu);
    g_h2d[g_h2d_count & 15].dst = d; g_h2d[g_h2d_count & 15].n = n; ++g_h2d_count;
    pthread_mutex_unlock(&g_mu);
  }
  return cudaSuccess;
}
API cudaError_t cudaMemcpy2D(void* d, size_t dp, const void* s, size_t sp, size_t w, size_t h, enum cudaMemcpyKind k) {
  for (size_t i = 0; i < h; ++i) memmove((char*)d + i * dp, (const char*)s + i * sp, w);
  return cudaSuccess;
}
static char g_symbol[1 << 20];
API cudaError_t cudaGetSymbolAddress(void** p, const void* sym) { GATE("cudaGetSymbolAddress"); *p = g_symbol; return cudaSuccess; }
API cudaError_t cudaMemcpyFromSymbolAsync(void* d, const void* sym, size_t n, size_t off, enum cudaMemcpyKind k, cudaStream_t s) { memset(d, 0, n); return cudaSuccess; }

API cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned f) { GATE("cudaStreamCreateWithFlags"); *s = (cudaStream_t)calloc(1, 8); return cudaSuccess; }
API cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return cudaSuccess; }
API cudaError_t cudaStreamSynchronize(cudaStream_t s) { GATE("cudaStreamSynchronize"); return cudaSuccess; }
API cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned f) { return cudaSuccess; }
API cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = (cudaEvent_t)calloc(1, 8); return cudaSuccess; }
API cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned f) { *e = (cudaEvent_t)calloc(1, 8); return cudaSuccess; }
API cudaError_t cudaEventDestroy(cudaEvent_t e) { free(e); return cudaSuccess; }
API cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s) { GATE("cudaEventRecord"); return cudaSuccess; }
API cudaError_t cudaEventQuery(cudaEvent_t e) { return cudaSuccess; }
API cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t a, cudaEvent_t b) { *ms = 0.f; return cudaSuccess; }

API cudaError_t cudaStreamBeginCapture(cudaStream_t s, enum cudaStreamCaptureMode m) { GATE("cudaStreamBeginCapture"); return cudaSuccess; }
API cudaError_t cudaStreamEndCapture(cudaStream_t s, cudaGraph_t* g) { *g = (cudaGraph_t)calloc(1, 8); return cudaSuccess; }
API cudaError_t cudaGraphInstantiate(cudaGraphExec_t* ge, cudaGraph_t g, unsigned long long flags) { *ge = (cudaGraphExec_t)calloc(1, 8); return cudaSuccess; }
API cudaError_t cudaGraphLaunch(cudaGraphExec_t ge, cudaStream_t s) { GATE("cudaGraphLaunch"); INC(g_graph_launches); return cudaSuccess; }
API cudaError_t cudaGraphDestroy(cudaGraph_t g) { free(g); return cudaSuccess; }
API cudaError_t cudaGraphExecDestroy(cudaGraphExec_t ge) { free(ge); return cudaSuccess; }

API cudaError_t cudaFuncSetAttribute(const void* f, enum cudaFuncAttribute a, int v) { return cudaSuccess; }
API cudaError_t cudaOccupancyMaxActiveClusters(int* n, const void* f, const cudaLaunchConfig_t* c) { *n = 14; return cudaSuccess; }
/* launch log: which kernel (device-side name from __cudaRegisterFunction), grid, block, dynamic shared memory */
#define MAX_FUNCS 256
#define MAX_LOG 8192
static struct { const void* host; const char* name; } g_funcs[MAX_FUNCS];
static int g_nfuncs = 0;
static struct { const char* name; unsigned gx, gy, gz, bx; size_t smem; } g_log[MAX_LOG];
static long g_nlog = 0;
API void fake_cuda_log_clear(void) { g_nlog = 0; }
API long fake_cuda_log_count(void) { return g_nlog < MAX_LOG ? g_nlog : MAX_LOG; }
API const char* fake_cuda_log_name(long i) { return g_log[i].name; }
API void fake_cuda_log_dims(long i, unsigned* out5, size_t* smem) {
  out5[0] = g_log[i].gx; out5[1] = g_log[i].gy; out5[2] = g_log[i].gz; out5[3] = g_log[i].bx; *smem = g_log[i].smem;
}
API cudaError_t cudaLaunchKernel(const void* f, dim3 g, dim3 b, void** args, size_t smem, cudaStream_t s) {
  GATE("cudaLaunchKernel");
  INC(g_launches);
  pthread_mutex_lock(&g_mu);
  if (g_nlog < MAX_LOG) {
    const char* name = "?";
    for (int i = 0; i < g_nfuncs; ++i) if (g_funcs[i].host == f) { name = g_funcs[i].name; break; }
    g_log[g_nlog].name = name; g_log[g_nlog].gx = g.x; g_log[g_nlog].gy = g.y; g_log[g_nlog].gz = g.z; g_log[g_nlog].bx = b.x;
    g_log[g_nlog].smem = smem;
  }
  ++g_nlog;
  pthread_mutex_unlock(&g_mu);
  return cudaSuccess;
}

static int fake_encode_tiled() { return 0; }          /* CUDA_SUCCESS: the tensor map stays zero, nothing reads it */
API cudaError_t cudaGetDriverEntryPoint(const char* sym, void** fn, unsigned long long flags, enum cudaDriverEntryPointQueryResult* q) {
  *fn = (void*)fake_encode_tiled;
  if (q) *q = cudaDriverEntryPointSuccess;
  return cudaSuccess;
}

/* what nvcc's host stubs call */
static __thread struct { dim3 g, b; size_t smem; void* st; } g_cfg;
API unsigned __cudaPushCallConfiguration(dim3 g, dim3 b, size_t smem, void* st) { g_cfg.g = g; g_cfg.b = b; g_cfg.smem = smem; g_cfg.st = st; return 0; }
API cudaError_t __cudaPopCallConfiguration(dim3* g, dim3* b, size_t* smem, void* st) { *g = g_cfg.g; *b = g_cfg.b; *smem = g_cfg.smem; *(void**)st = g_cfg.st; return cudaSuccess; }
static void* g_fatbin_handle;
API void** __cudaRegisterFatBinary(void* fatbin) { return &g_fatbin_handle; }
API void __cudaRegisterFatBinaryEnd(void** h) {}
API void __cudaUnregisterFatBinary(void** h) {}
API void __cudaRegisterFunction(void** h, const char* hf, char* df, const char* dn, int tl, void* tid, void* bid, void* bd, void* gd, int* ws) {
  if (g_nfuncs < MAX_FUNCS) { g_funcs[g_nfuncs].host = hf; g_funcs[g_nfuncs].name = dn; ++g_nfuncs; }
}
API void __cudaRegisterVar(void** h, char* hv, char* da, const char* dn, int ext, size_t size, int constant, int global) {}
