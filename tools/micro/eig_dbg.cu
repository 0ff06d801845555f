// host vs device check of csrc/eeg_eig6.cuh on one matrix (nvcc -I energy_equiv_lattice_gnn_b200/csrc)
#include <cstdio>
#include <cuda_runtime.h>
#include "eeg_eig6.cuh"
__global__ void k(const float* m, float* ev) {
  __shared__ float sm[36];
  for (int t = threadIdx.x; t < 36; t += 32) sm[t] = m[t];
  __syncwarp();
  if (threadIdx.x == 1) eeg_eigvalsh6(sm, ev);
  if (threadIdx.x == 0) eeg_eigvalsh6(m, ev + 6);
}
int main() {
  float m[36];
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) m[i * 6 + j] = 1e-3f * (1.f / (1 + i + j) + (i == j ? 0.3f * i : 0));
  float ev[6];
  eeg_eigvalsh6(m, ev);
  printf("host: ");
  for (int i = 0; i < 6; i++) printf("%g ", ev[i]);
  float *dm, *de;
  cudaMalloc(&dm, 144);
  cudaMalloc(&de, 48);
  cudaMemcpy(dm, m, 144, cudaMemcpyHostToDevice);
  k<<<1, 32>>>(dm, de);
  float h[12];
  cudaMemcpy(h, de, 48, cudaMemcpyDeviceToHost);
  printf("\ndev smem: ");
  for (int i = 0; i < 6; i++) printf("%g ", h[i]);
  printf("\ndev gmem: ");
  for (int i = 6; i < 12; i++) printf("%g ", h[i]);
  printf("\n%s\n", cudaGetErrorString(cudaGetLastError()));
}
