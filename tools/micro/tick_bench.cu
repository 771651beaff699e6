// Is the cost of a kernel in a stream quantised?  Spin kernels of controlled length, back to back, plain launches, PDL, graph.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void spin(long long cycles, int pdl) {
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {}
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
int main() {
  cudaStream_t st; cudaStreamCreate(&st);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int N = 400;
  for (int mode = 0; mode < 3; ++mode) {
    for (long long cyc : {0LL, 1000LL, 2000LL, 3000LL, 4000LL, 5000LL, 6000LL, 8000LL, 10000LL, 12000LL, 16000LL}) {
      float ms = 0;
      if (mode == 0) {
        for (int i = 0; i < 20; ++i) spin<<<148, 128, 0, st>>>(cyc, 0);
        cudaEventRecord(e0, st);
        for (int i = 0; i < N; ++i) spin<<<148, 128, 0, st>>>(cyc, 0);
        cudaEventRecord(e1, st); cudaEventSynchronize(e1);
      } else if (mode == 1) {
        cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.stream = st;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        for (int i = 0; i < 20; ++i) cudaLaunchKernelEx(&cfg, spin, cyc, 1);
        cudaEventRecord(e0, st);
        for (int i = 0; i < N; ++i) cudaLaunchKernelEx(&cfg, spin, cyc, 1);
        cudaEventRecord(e1, st); cudaEventSynchronize(e1);
      } else {
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < N; ++i) spin<<<148, 128, 0, st>>>(cyc, 0);
        cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, st); cudaStreamSynchronize(st);
        cudaEventRecord(e0, st); cudaGraphLaunch(ge, st); cudaEventRecord(e1, st); cudaEventSynchronize(e1);
        cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
      }
      cudaEventElapsedTime(&ms, e0, e1);
      printf("%s spin %lld cycles: %.3f us per kernel (%s)\n", mode == 0 ? "plain" : mode == 1 ? "pdl" : "graph", cyc, ms * 1000 / N, cudaGetErrorString(cudaGetLastError()));
    }
  }
  return 0;
}
