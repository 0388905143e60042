// copy_probe.cu -- bare host<->device copy ceiling of this box for the bench's e2e traffic pattern (VERDICT r1 #5a).
// For n = 1, 2, 4, 8 (as many as there are GPUs): n host threads, one GPU each, every thread loops
//   cudaMemcpyAsync H2D of 1.44 GB pinned (one step's images) on one stream, D2H of 0.36 GB pinned (one step's result
//   slabs) on a second stream, both in flight together, `reps` times.
// Prints one JSON line per n: aggregate GB/s for H2D alone, D2H alone and both together.  No kernels run.
// Build: nvcc -O2 -o copy_probe copy_probe.cu -lpthread
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#define CK(x)                                                                            \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) {                                                             \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
      exit(2);                                                                           \
    }                                                                                    \
  } while (0)

static const size_t kH2D = 2000ull * 2 * 752 * 480;               // 1.444 GB
static const size_t kD2H = 2000ull * (2 * 1500 * 56 + 1500 * 8);  // 0.36 GB

struct Worker {
  int dev;
  void *hin = nullptr, *hout = nullptr, *din = nullptr, *dout = nullptr;
  cudaStream_t s0, s1;
};

static std::atomic<int> g_arrived;
static void barrier(int target) {
  g_arrived.fetch_add(1);
  while (g_arrived.load() < target) std::this_thread::yield();
}

int main(int argc, char** argv) {
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const int reps = argc > 1 ? atoi(argv[1]) : 5;
  std::vector<Worker> w(ndev);
  for (int d = 0; d < ndev; ++d) {
    w[d].dev = d;
    CK(cudaSetDevice(d));
    CK(cudaMallocHost(&w[d].hin, kH2D));
    CK(cudaMallocHost(&w[d].hout, kD2H));
    memset(w[d].hin, d + 1, kH2D);
    memset(w[d].hout, 0, kD2H);
    CK(cudaMalloc(&w[d].din, kH2D));
    CK(cudaMalloc(&w[d].dout, kD2H));
    CK(cudaStreamCreate(&w[d].s0));
    CK(cudaStreamCreate(&w[d].s1));
  }
  for (int n = 1; n <= ndev; n *= 2) {
    double secs[3] = {0, 0, 0};
    for (int mode = 0; mode < 3; ++mode) {  // 0: H2D, 1: D2H, 2: both
      g_arrived = 0;
      std::vector<std::thread> th;
      std::vector<double> t(n);
      for (int d = 0; d < n; ++d)
        th.emplace_back([&, d]() {
          CK(cudaSetDevice(d));
          for (int it = 0; it < 2; ++it) {  // warm-up
            if (mode != 1) CK(cudaMemcpyAsync(w[d].din, w[d].hin, kH2D, cudaMemcpyHostToDevice, w[d].s0));
            if (mode != 0) CK(cudaMemcpyAsync(w[d].hout, w[d].dout, kD2H, cudaMemcpyDeviceToHost, w[d].s1));
          }
          CK(cudaDeviceSynchronize());
          barrier(n);
          const auto t0 = std::chrono::steady_clock::now();
          for (int it = 0; it < reps; ++it) {
            if (mode != 1) CK(cudaMemcpyAsync(w[d].din, w[d].hin, kH2D, cudaMemcpyHostToDevice, w[d].s0));
            if (mode != 0) CK(cudaMemcpyAsync(w[d].hout, w[d].dout, kD2H, cudaMemcpyDeviceToHost, w[d].s1));
          }
          CK(cudaDeviceSynchronize());
          barrier(2 * n);  // everybody done: the slowest GPU defines the time
          t[d] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        });
      for (auto& x : th) x.join();
      for (double x : t) secs[mode] = x > secs[mode] ? x : secs[mode];
    }
    const double gb_in = (double)kH2D * reps * n / 1e9, gb_out = (double)kD2H * reps * n / 1e9;
    printf("{\"probe\": \"copy\", \"gpus\": %d, \"reps\": %d, \"h2d_bytes\": %zu, \"d2h_bytes\": %zu, "
           "\"h2d_alone_gbs\": %.1f, \"d2h_alone_gbs\": %.1f, \"both_h2d_gbs\": %.1f, \"both_total_gbs\": %.1f, "
           "\"both_ms_per_step\": %.2f, \"pairs_per_s_ceiling\": %.0f}\n",
           n, reps, kH2D, kD2H, gb_in / secs[0], gb_out / secs[1], gb_in / secs[2], (gb_in + gb_out) / secs[2],
           1e3 * secs[2] / reps, 2000.0 * n * reps / secs[2]);
    fflush(stdout);
  }
  return 0;
}
