#include <cstdio>
#include <cstdint>
__device__ __forceinline__ int ring_offset(int k, int pitch) {
  const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  return dx[k] + dy[k] * pitch;
}
__device__ __forceinline__ int pmin(int a, int b) { int r; asm("min.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ int pmax(int a, int b) { int r; asm("max.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ int smin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int smax(int a, int b) { return a > b ? a : b; }
template <int V>
__device__ int score(const uint8_t* p, int pitch) {
  const int v = p[0];
  int d[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) d[k] = v - (int)p[ring_offset(k, pitch)];
  int best = -1000;
#pragma unroll
  for (int s = 0; s < 16; ++s) {
    int mn = d[s], mx = d[s];
#pragma unroll
    for (int i = 1; i < 9; ++i) {
      const int x = d[(s + i) & 15];
      if (V == 0) { mn = min(mn, x); mx = max(mx, x); }
      if (V == 1) { mn = pmin(mn, x); mx = pmax(mx, x); }
      if (V == 2) { mn = smin(mn, x); mx = smax(mx, x); }
    }
    if (V == 0) best = max(best, max(mn, -mx));
    if (V == 1) best = pmax(best, pmax(mn, -mx));
    if (V == 2) best = smax(best, smax(mn, -mx));
    if (V == 3) {}
  }
  return best - 1;
}
// variant 3: only the dark side / only the bright side kept apart
__device__ int score_split(const uint8_t* p, int pitch) {
  const int v = p[0];
  int d[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) d[k] = v - (int)p[ring_offset(k, pitch)];
  int a = -1000, b = 1000;
#pragma unroll
  for (int s = 0; s < 16; ++s) {
    int mn = d[s], mx = d[s];
#pragma unroll
    for (int i = 1; i < 9; ++i) { mn = min(mn, d[(s + i) & 15]); mx = max(mx, d[(s + i) & 15]); }
    a = max(a, mn);
    b = min(b, mx);
  }
  return max(a, -b) - 1;
}
__global__ void tk(const uint8_t* img, int* out) {
  __shared__ uint8_t s[7][80];
  for (int i = threadIdx.x; i < 49; i += blockDim.x) s[i / 7][i % 7] = img[i];
  __syncthreads();
  if (threadIdx.x == 0) {
    out[0] = score<0>(&s[3][3], 80);
    out[1] = score<1>(&s[3][3], 80);
    out[2] = score<2>(&s[3][3], 80);
    out[3] = score_split(&s[3][3], 80);
  }
}
int main() {
  uint8_t img[7][7]; int v = 183; int d[16] = {106,105,65,6,-7,25,52,57,68,95,134,164,157,147,110,109};
  const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1}; const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  for (int y = 0; y < 7; y++) for (int x = 0; x < 7; x++) img[y][x] = v;
  for (int k = 0; k < 16; k++) img[3 + dy[k]][3 + dx[k]] = v - d[k];
  uint8_t* di; int* dout; cudaMalloc(&di, 49); cudaMalloc(&dout, 100 * 4);
  cudaMemcpy(di, img, 49, cudaMemcpyHostToDevice);
  tk<<<1, 32>>>(di, dout);
  int out[32]; cudaMemcpy(out, dout, 4 * 4, cudaMemcpyDeviceToHost);
  printf("builtin %d  ptx %d  select %d  split %d  (want 94)\n", out[0], out[1], out[2], out[3]);
  return 0;
}
