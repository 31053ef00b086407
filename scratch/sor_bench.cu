// micro-benchmark of the SOR kernels (scratch, not part of the product)
#include <cstdio>
#include <vector>
#include <random>
#include <cstring>
#define SOR_PROFILE
#include "../of_dis_b200/csrc/varref.cu"
using namespace disb200;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__);return 1;}}while(0)
int main(int argc,char**argv){
  int sizes[3][2]={{32,14},{64,28},{128,56}};
  int iters=3; float omega=1.6f;
  for(int np: {256,1024}) for(auto&sz:sizes){
    int w=sz[0],h=sz[1]; size_t plane=((size_t)w*h+3)&~3ull;
    std::mt19937 rng(1); std::uniform_real_distribution<float> U(0.5f,5.f); std::normal_distribution<float> N(0,1);
    std::vector<float4> A(plane*np),B(plane*np);
    for(int p=0;p<np;++p) for(int j=0;j<h;++j) for(int i=0;i<w;++i){
      size_t k=p*plane+j*w+i; float hh=i<w-1?U(rng):0.f, vv=j<h-1?U(rng):0.f;
      A[k]=make_float4(hh,vv,0.f,0.01f*N(rng)); B[k]=make_float4(0.02f+0.004f*U(rng),0.02f+0.004f*U(rng),N(rng),N(rng));
    }
    for(int p=0;p<np;++p) for(int j=1;j<h;++j) for(int i=0;i<w;++i) A[p*plane+j*w+i].z=A[p*plane+(j-1)*w+i].y;
    float4 *dA,*dB; float *du1,*du2;
    CK(cudaMalloc(&dA,sizeof(float4)*plane*np)); CK(cudaMalloc(&dB,sizeof(float4)*plane*np));
    CK(cudaMalloc(&du1,sizeof(float)*2*plane*np)); CK(cudaMalloc(&du2,sizeof(float)*2*plane*np));
    CK(cudaMemcpy(dA,A.data(),sizeof(float4)*plane*np,cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB,B.data(),sizeof(float4)*plane*np,cudaMemcpyHostToDevice));
    Dims d{w,h,8,w+16,0,plane};
    int threads=((h*iters+31)/32)*32;
    size_t smem=((size_t)(h+2)*w+4)*sizeof(float2);
    CK(cudaFuncSetAttribute(sor_lean_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms1=0,ms2=0,ms3=0; const int reps=20;
    float* du3; CK(cudaMalloc(&du3,sizeof(float)*2*plane*np));
    const int G=(h+31)/32; const int rthreads=32*iters*G<128?128:32*iters*G; const size_t rsmem=((sorr_du_elems(w,h)*2+3)&~3ull)*4+sorr_ring_bytes(h,3);
    CK(cudaFuncSetAttribute(sor_rows_kernel<true,3>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    for(int v=0;v<3;++v){
      float* du=v==0?du1:(v==1?du2:du3);
      for(int rep=-3;rep<reps;++rep){
        if(rep<=0){ CK(cudaMemset(du,0,sizeof(float)*2*plane*np)); }
        if(rep==0) cudaEventRecord(e0);
        if(v==0) sor_wavefront_kernel<true><<<np,threads>>>(du,dA,dB,d,iters,omega,1);
        else if(v==1) sor_lean_kernel<<<np,threads,smem>>>(du,dA,dB,d,iters,omega);
        else sor_rows_kernel<true,3><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega);
      }
      cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); float ms; cudaEventElapsedTime(&ms,e0,e1);
      if(v==1){ long long pr[8]; cudaMemcpyFromSymbol(pr,g_sor_prof,sizeof(pr)); printf("   v%d: prologue-sync %lld  loop %lld cycles (%.0f per step)  epilogue %lld\n",v,pr[1]-pr[0],pr[2]-pr[1],(double)(pr[2]-pr[1])/(v==1?(w/4+h-1+4):(w+ (h<32?h:32) -1)),pr[3]-pr[2]); } (v==0?ms1:(v==1?ms2:ms3))=ms/reps;
      CK(cudaGetLastError());
    }
    // compare after identical number of calls
    std::vector<float> r1(2*plane*np), r2(2*plane*np);
    CK(cudaMemcpy(r1.data(),du1,r1.size()*4,cudaMemcpyDeviceToHost)); CK(cudaMemcpy(r2.data(),du2,r2.size()*4,cudaMemcpyDeviceToHost));
    std::vector<float> r3(2*plane*np); CK(cudaMemcpy(r3.data(),du3,r3.size()*4,cudaMemcpyDeviceToHost));
    size_t nd=0; double md=0; for(size_t k=0;k<r1.size();++k){ if(r1[k]!=r2[k]){++nd; md=fmax(md,fabs(r1[k]-r2[k]));} }
    size_t nd3=0; for(size_t k=0;k<r1.size();++k) if(r1[k]!=r3[k]) ++nd3;
    int nchunks=w/4; int steps=nchunks+h-1+2*(iters-1);
    printf("pairs %4d  %3dx%2d: general %.1f us  lean %.1f us (%d macro-steps, %.0f cycles/step @1.9GHz) rows %.1f us (%d steps, %.0f cycles/step)  mismatches lean %zu rows %zu maxdiff %.3g  sample %.5f\n",np,w,h,ms1*1e3,ms2*1e3,steps,ms2*1e-3*1.9e9/steps,ms3*1e3,w+h-1+4,ms3*1e-3*1.9e9/(w+h-1+4),nd,nd3,md,r2[plane]);
    cudaFree(dA);cudaFree(dB);cudaFree(du1);cudaFree(du2);cudaFree(du3);
  }
  return 0;
}
