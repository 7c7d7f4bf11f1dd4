// dev probe: do cooperative launches overlap (a) each other, (b) ordinary kernels on other streams?
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;
__global__ void __launch_bounds__(256, 2) spin_coop(long long ns, int syncs) {
  cg::grid_group g = cg::this_grid();
  long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (int s = 0; s < syncs; s++) {
    long long t; do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns * (s + 1) / syncs);
    g.sync();
  }
}
__global__ void __launch_bounds__(256, 2) spin_plain(long long ns) {
  long long t0; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  long long t; do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns);
}
static float run(int mode, int ctas) {
  cudaStream_t a, b; cudaStreamCreateWithFlags(&a, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&b, cudaStreamNonBlocking);
  cudaEvent_t e0, e1, ea, eb; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&ea); cudaEventCreate(&eb);
  long long ns = 200000; int syncs = 5; void* args[] = {&ns, &syncs};
  float best = 1e9;
  for (int rep = 0; rep < 5; rep++) {
    cudaDeviceSynchronize();
    cudaEventRecord(e0, a); cudaStreamWaitEvent(b, e0, 0);
    if (mode == 0) { cudaLaunchCooperativeKernel((void*)spin_coop, dim3(ctas), dim3(256), args, 0, a); cudaLaunchCooperativeKernel((void*)spin_coop, dim3(ctas), dim3(256), args, 0, b); }
    if (mode == 1) { cudaLaunchCooperativeKernel((void*)spin_coop, dim3(ctas), dim3(256), args, 0, a); spin_plain<<<ctas, 256, 0, b>>>(ns); }
    if (mode == 2) { spin_plain<<<ctas, 256, 0, a>>>(ns); spin_plain<<<ctas, 256, 0, b>>>(ns); }
    if (mode == 3) { spin_plain<<<ctas, 256, 0, a>>>(ns); cudaLaunchCooperativeKernel((void*)spin_coop, dim3(ctas), dim3(256), args, 0, b); }
    cudaEventRecord(eb, b); cudaStreamWaitEvent(a, eb, 0); cudaEventRecord(e1, a);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best * 1e3f;
}
int main() {
  const char* names[] = {"coop + coop", "coop then plain", "plain + plain", "plain then coop"};
  for (int ctas : {37, 74, 148})
    for (int m = 0; m < 4; m++) printf("ctas %3d  %-16s two 200 us kernels on two streams: %.0f us\n", ctas, names[m], run(m, ctas));
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
