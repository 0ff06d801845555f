#include <cmath>
#include <cstdio>
#include <cuda_runtime.h>
__host__ __device__ void eigvalsh6(const float* __restrict__ src, float* __restrict__ ev) {
  float a[6][6];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j <= i; ++j) a[i][j] = a[j][i] = src[i * 6 + j];  // lower triangle, like LAPACK 'L' behind eigvalsh
  for (int sweep = 0; sweep < 12; ++sweep) {
    float off = 0.f, diag = 0.f;
    for (int i = 0; i < 6; ++i) {
      diag += a[i][i] * a[i][i];
      for (int j = 0; j < i; ++j) off += a[i][j] * a[i][j];
    }
    if (off <= 1e-16f * diag) break;
    for (int p = 0; p < 5; ++p)
      for (int q = p + 1; q < 6; ++q) {
        const float apq = a[p][q];
        if (apq == 0.f) continue;
        const float theta = (a[q][q] - a[p][p]) / (2.f * apq);
        const float t = (theta >= 0.f ? 1.f : -1.f) / (fabsf(theta) + sqrtf(fmaf(theta, theta, 1.f)));
        const float c = rsqrtf(fmaf(t, t, 1.f)), s = t * c;
        for (int k = 0; k < 6; ++k) {
          const float akp = a[k][p], akq = a[k][q];
          a[k][p] = c * akp - s * akq;
          a[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 6; ++k) {
          const float apk = a[p][k], aqk = a[q][k];
          a[p][k] = c * apk - s * aqk;
          a[q][k] = s * apk + c * aqk;
        }
      }
  }
  float e[6];
  for (int i = 0; i < 6; ++i) e[i] = a[i][i];
  for (int i = 1; i < 6; ++i) {
    const float v = e[i];
    int j = i - 1;
    while (j >= 0 && e[j] > v) {
      e[j + 1] = e[j];
      --j;
    }
    e[j + 1] = v;
  }
  for (int i = 0; i < 6; ++i) ev[i] = e[i];
}


__global__ void k(const float* m, float* ev){ __shared__ float sm[36]; if(threadIdx.x<32){ for(int t=threadIdx.x;t<36;t+=32) sm[t]=m[t]; } __syncwarp(); if(threadIdx.x==1) eigvalsh6(sm,ev); if(threadIdx.x==0) eigvalsh6(m,ev+6);}
int main(){ float m[36]; for(int i=0;i<6;i++)for(int j=0;j<6;j++){ m[i*6+j]= 1e-3f*(1.f/(1+i+j) + (i==j? 0.3f*i:0));} float ev[6]; eigvalsh6(m,ev); printf("host: "); for(int i=0;i<6;i++)printf("%g ",ev[i]); printf("\n");
 float *dm,*de; cudaMalloc(&dm,144); cudaMalloc(&de,48); cudaMemcpy(dm,m,144,cudaMemcpyHostToDevice); k<<<1,32>>>(dm,de); float h[12]; cudaMemcpy(h,de,48,cudaMemcpyDeviceToHost); printf("dev smem: "); for(int i=0;i<6;i++)printf("%g ",h[i]); printf("\ndev gmem: "); for(int i=6;i<12;i++)printf("%g ",h[i]); printf("\n%s\n", cudaGetErrorString(cudaGetLastError())); }