// Microbenchmark: FP64 DFMA vs DMMA.8x8x4 issue-bound peak on sm_100a, plus fragment layout check.
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include <cmath>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1;}}while(0)

template<int NACC>
__global__ void dfma_kernel(double* out, int iters, double a, double b){
  double acc[NACC];
  #pragma unroll
  for(int i=0;i<NACC;i++) acc[i]=threadIdx.x*1e-9+i;
  for(int it=0; it<iters; it++){
    #pragma unroll
    for(int i=0;i<NACC;i++) acc[i]=fma(acc[i],a,b);
  }
  double s=0;
  #pragma unroll
  for(int i=0;i<NACC;i++) s+=acc[i];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

template<int NT>
__global__ void dmma_kernel(double* out, int iters, double a0, double b0){
  double c[NT][2];
  #pragma unroll
  for(int i=0;i<NT;i++){c[i][0]=threadIdx.x*1e-9; c[i][1]=i;}
  double a=a0+threadIdx.x*1e-12, b=b0;
  for(int it=0; it<iters; it++){
    #pragma unroll
    for(int i=0;i<NT;i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[i][0]),"+d"(c[i][1]) : "d"(a),"d"(b));
  }
  double s=0;
  #pragma unroll
  for(int i=0;i<NT;i++) s+=c[i][0]+c[i][1];
  out[blockIdx.x*blockDim.x+threadIdx.x]=s;
}

// layout check: C(8x8) = A(8x4) * B(4x8)
__global__ void layout884(const double* A, const double* B, double* C){
  int lane=threadIdx.x, g=lane>>2, t=lane&3;
  double a=A[g*4+t], b=B[t*8+g]; double c0=0,c1=0;
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0),"+d"(c1) : "d"(a),"d"(b));
  C[g*8+2*t]=c0; C[g*8+2*t+1]=c1;
}
// C(16x8) = A(16x16)*B(16x8)
__global__ void layout16816(const double* A, const double* B, double* C){
  int lane=threadIdx.x, g=lane>>2, t=lane&3;
  double av[8], bv[4];
  for(int i=0;i<4;i++){ av[2*i]=A[g*16+t+4*i]; av[2*i+1]=A[(g+8)*16+t+4*i]; bv[i]=B[(t+4*i)*8+g]; }
  double c0=0,c1=0,c2=0,c3=0;
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};" : "+d"(c0),"+d"(c1),"+d"(c2),"+d"(c3) : "d"(av[0]),"d"(av[1]),"d"(av[2]),"d"(av[3]),"d"(av[4]),"d"(av[5]),"d"(av[6]),"d"(av[7]),"d"(bv[0]),"d"(bv[1]),"d"(bv[2]),"d"(bv[3]));
  C[g*8+2*t]=c0; C[g*8+2*t+1]=c1; C[(g+8)*8+2*t]=c2; C[(g+8)*8+2*t+1]=c3;
}

int main(){
  int dev=0; cudaDeviceProp p; CK(cudaGetDeviceProperties(&p,dev));
  int clk=0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  printf("dev %s SMs %d clockRate %d kHz\n", p.name, p.multiProcessorCount, clk);
  double* out; CK(cudaMalloc(&out, sizeof(double)*148*64*1024));
  cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int nsm=p.multiProcessorCount;
  for(int wpb : {4,8,16,32}){
    for(int bps: {1,2}){
      if(wpb*bps>64) continue;
      int iters=20000; float ms;
      dfma_kernel<8><<<nsm*bps, wpb*32>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      cudaEventRecord(e0); dfma_kernel<8><<<nsm*bps, wpb*32>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      cudaEventElapsedTime(&ms,e0,e1);
      double fl=2.0*8*iters*(double)nsm*bps*wpb*32;
      printf("DFMA  warps/blk %2d blk/SM %d : %.2f TFLOP/s (%.3f ms)\n", wpb,bps, fl/ms*1e-9, ms);
      dmma_kernel<8><<<nsm*bps, wpb*32>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      cudaEventRecord(e0); dmma_kernel<8><<<nsm*bps, wpb*32>>>(out, iters, 1.0000001, 1e-9); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
      cudaEventElapsedTime(&ms,e0,e1);
      fl=2.0*256*8*iters*(double)nsm*bps*wpb;
      printf("DMMA  warps/blk %2d blk/SM %d : %.2f TFLOP/s (%.3f ms)\n", wpb,bps, fl/ms*1e-9, ms);
    }
  }
  // layout checks
  std::vector<double> A(256),B(128),C(128),R(128);
  for(int i=0;i<256;i++)A[i]=sin(i*0.37)+0.1*i; for(int i=0;i<128;i++)B[i]=cos(i*0.11)-0.01*i;
  double *dA,*dB,*dC; cudaMalloc(&dA,256*8);cudaMalloc(&dB,128*8);cudaMalloc(&dC,128*8);
  cudaMemcpy(dA,A.data(),256*8,cudaMemcpyHostToDevice);cudaMemcpy(dB,B.data(),128*8,cudaMemcpyHostToDevice);
  layout884<<<1,32>>>(dA,dB,dC); CK(cudaDeviceSynchronize()); cudaMemcpy(C.data(),dC,64*8,cudaMemcpyDeviceToHost);
  double err=0; for(int i=0;i<8;i++)for(int j=0;j<8;j++){double s=0;for(int k=0;k<4;k++)s+=A[i*4+k]*B[k*8+j]; err=fmax(err,fabs(s-C[i*8+j]));}
  printf("layout m8n8k4 max err %.3e\n",err);
  layout16816<<<1,32>>>(dA,dB,dC); CK(cudaDeviceSynchronize()); cudaMemcpy(C.data(),dC,128*8,cudaMemcpyDeviceToHost);
  err=0; for(int i=0;i<16;i++)for(int j=0;j<8;j++){double s=0;for(int k=0;k<16;k++)s+=A[i*16+k]*B[k*8+j]; err=fmax(err,fabs(s-C[i*8+j]));}
  printf("layout m16n8k16 max err %.3e\n",err);
  return 0;
}
