// scratch: variants of the row-per-lane SOR, timed against each other (not part of the product)
#include <cstdio>
#include <vector>
#include <random>
#include <cstring>
#include "../of_dis_b200/csrc/varref.cu"
using namespace disb200;
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("CUDA %s at %d\n",cudaGetErrorString(e),__LINE__);return 1;}}while(0)

__device__ long long g_prof[8];

// V1: hand-scheduled loop.  FLAGS bit0: no named barrier (timing only), bit1: no cp.async (reads stale ring)
template <int NS, int FLAGS>
__global__ void __launch_bounds__(512)
sorr_v1(float* __restrict__ du_all, const float4* __restrict__ cA_all, const float4* __restrict__ cB_all, Dims d, float omega) {
  extern __shared__ float4 smem[];
  const int pair = blockIdx.x, w = d.w, h = d.h, tid = threadIdx.x, NT = blockDim.x;
  float* sdu = reinterpret_cast<float*>(smem);
  const size_t du_floats = (sorr_du_elems(w, h) * 2 + 3) & ~(size_t)3;
  float4* ring = reinterpret_cast<float4*>(sdu + du_floats);
  float* gdu = du_all + (size_t)pair * d.plane * 2;
  const int npix = w * h;
  for (int k = tid; k < w * 2; k += NT) { sdu[k] = 0.f; sdu[(size_t)(h + 1) * w * 2 + k] = 0.f; }
  for (int k = tid; k < 4; k += NT) sdu[(size_t)(h + 2) * w * 2 + k] = 0.f;
  for (int k = tid; k < npix * 2; k += NT) sdu[(size_t)w * 2 + k] = gdu[k];
  __syncthreads();
  const int nb = (h + 31) >> 5;
  const int nwarps = NS * nb;
  if ((tid >> 5) < nwarps) {
    constexpr int RING = 16;   // D + 2(NS-1) + 1 <= 16
    constexpr int D = SORR_D;
    const int warp = tid >> 5, lane = tid & 31;
    const int s = warp / nb, b = warp - s * nb;
    const int j = b * 32 + lane;
    const bool row = j < h;
    const bool mid = j > 0 && j < h - 1, first = j == 0;
    const int nthreads = nwarps * 32;
    const int T = w + h - 1 + 2 * (NS - 1);
    const float4* cA = cA_all + (size_t)pair * d.plane;
    const float4* cB = cB_all + (size_t)pair * d.plane;
    char* rbase = reinterpret_cast<char*>(ring + (size_t)b * RING * 64) + lane * 16;   // cA slots; cB at + RING*512
    const unsigned uw = (unsigned)w;
    int i = row ? -j - 2 * s : -(1 << 28);              // pixel of this thread at step t = 0
    float2* p = reinterpret_cast<float2*>(sdu) + (size_t)(row ? j + 1 : 1) * w + (row ? i : 0);  // &du[j][i]
    unsigned roff = ((unsigned)i & (RING - 1)) * 512u;
    // coefficient stream
    const float4* gA = cA + (size_t)(row ? j : 0) * w + (row ? i : 0);
    const float4* gB = cB + (size_t)(row ? j : 0) * w + (row ? i : 0);
    if (s == 0 && !(FLAGS & 2)) {
      for (int k = 0; k < D; ++k) {
        const int ip = i + k;
        if ((unsigned)ip < uw) {
          cp_async16(rbase + (((unsigned)ip & (RING - 1)) * 512u), gA + k);
          cp_async16(rbase + (((unsigned)ip & (RING - 1)) * 512u) + RING * 512, gB + k);
        }
        cp_async_commit();
      }
    }
    unsigned poff = ((unsigned)(i + D) & (RING - 1)) * 512u;
    gA += D; gB += D;
    float2 left = make_float2(0.f, 0.f), cur = left;
    float hprev = 0.f;
    if (blockIdx.x == 0 && tid == 0) g_prof[0] = clock64();
#pragma unroll 2
    for (int t = 0; t < T; ++t) {
      if (s == 0 && !(FLAGS & 2)) {
        if ((unsigned)(i + D) < uw) {
          cp_async16(rbase + poff, gA);
          cp_async16(rbase + poff + RING * 512, gB);
        }
        cp_async_commit();
        cp_async_wait<D>();
      }
      if ((unsigned)(i + 1) <= uw) {                    // i in [-1, w-1]
        const float2 r = p[1];
        if (i >= 0) {
          const float4 a = *reinterpret_cast<const float4*>(rbase + roff);
          const float4 c = *reinterpret_cast<const float4*>(rbase + roff + RING * 512);
          const float2 tp = p[-w], bt = p[w];
          const float2 nw = sorr_update_of(a, c, hprev, mid, first, omega, left, cur, r, tp, bt);
          p[0] = nw;
          left = nw;
          hprev = a.x;
        }
        cur = r;
      }
      ++i; ++p; ++gA; ++gB;
      roff = (roff + 512u) & (RING * 512u - 1);
      poff = (poff + 512u) & (RING * 512u - 1);
      if (FLAGS & 1) __syncwarp();
      else if (nthreads == 32) __syncwarp();
      else named_barrier(1, nthreads);
    }
    if (blockIdx.x == 0 && tid == 0) g_prof[1] = clock64();
  }
  __syncthreads();
  for (int k = tid; k < npix * 2; k += NT) gdu[k] = sdu[(size_t)w * 2 + k];
}

int main(int argc,char**argv){
  int sizes[3][2]={{32,14},{64,28},{128,56}};
  const int iters=3; float omega=1.6f;
  int np_sel=argc>1?atoi(argv[1]):0; int vmask=argc>2?atoi(argv[2]):0x7f;
  for(int np: {256,1184,2368}) for(auto&sz:sizes){ if(np_sel&&np!=np_sel) continue;
    int w=sz[0],h=sz[1]; size_t plane=((size_t)w*h+3)&~3ull;
    std::mt19937 rng(1); std::uniform_real_distribution<float> U(0.5f,5.f); std::normal_distribution<float> N(0,1);
    std::vector<float4> A(plane*np),B(plane*np);
    for(int p=0;p<np;++p) for(int j=0;j<h;++j) for(int i=0;i<w;++i){
      size_t k=p*plane+j*w+i; float hh=i<w-1?U(rng):0.f, vv=j<h-1?U(rng):0.f;
      A[k]=make_float4(hh,vv,0.f,0.01f*N(rng)); B[k]=make_float4(0.02f+0.004f*U(rng),0.02f+0.004f*U(rng),N(rng),N(rng));
    }
    for(int p=0;p<np;++p) for(int j=1;j<h;++j) for(int i=0;i<w;++i) A[p*plane+j*w+i].z=A[p*plane+(j-1)*w+i].y;
    float4 *dA,*dB;
    CK(cudaMalloc(&dA,sizeof(float4)*plane*np)); CK(cudaMalloc(&dB,sizeof(float4)*plane*np));
    CK(cudaMemcpy(dA,A.data(),sizeof(float4)*plane*np,cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB,B.data(),sizeof(float4)*plane*np,cudaMemcpyHostToDevice));
    Dims d{w,h,8,w+16,0,plane};
    const int G=(h+31)/32; const int rthreads=32*iters*G<128?128:32*iters*G;
    const size_t rsmem=((sorr_du_elems(w,h)*2+3)&~3ull)*4+sorr_ring_bytes(h,3);
    int threads=((h*iters+31)/32)*32;
    size_t lsmem=((size_t)(h+2)*w+4)*sizeof(float2);
    CK(cudaFuncSetAttribute(sor_lean_kernel,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sor_rows_kernel<true,3>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sorr_v1<3,0>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sorr_v1<3,1>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sorr_v1<3,2>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sorr_v1<3,3>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    const size_t wsmem=sorw_smem_bytes(h,3);
    CK(cudaFuncSetAttribute(sor_warp_kernel<true,3,1>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    CK(cudaFuncSetAttribute(sor_warp_kernel<true,3,2>,cudaFuncAttributeMaxDynamicSharedMemorySize,200*1024));
    cudaEvent_t e0,e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int NV=7; const int reps=20;
    std::vector<std::vector<float>> res(NV);
    float ms[NV]; long long cyc[NV];
    for(int v=0;v<NV;++v){ if(!((vmask>>v)&1)&&v!=0){ms[v]=0;cyc[v]=0;res[v]=res[0];continue;}
      float* du; CK(cudaMalloc(&du,sizeof(float)*2*plane*np));
      for(int rep=-3;rep<reps;++rep){
        if(rep<=0){ CK(cudaMemset(du,0,sizeof(float)*2*plane*np)); }
        if(rep==0) cudaEventRecord(e0);
        switch(v){
          case 0: sor_lean_kernel<<<np,threads,lsmem>>>(du,dA,dB,d,iters,omega); break;
          case 1: sor_rows_kernel<true,3><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega); break;
          case 2: sorr_v1<3,0><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega); break;
          case 3: sorr_v1<3,1><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega); break;
          case 4: sorr_v1<3,2><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega); break;
          case 5: sorr_v1<3,3><<<np,rthreads,rsmem>>>(du,dA,dB,d,omega); break;
          case 6: if(h<=32) sor_warp_kernel<true,3,1><<<np,32,wsmem>>>(du,dA,dB,d,omega); else sor_warp_kernel<true,3,2><<<np,32,wsmem>>>(du,dA,dB,d,omega); break;
        }
      }
      cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); cudaEventElapsedTime(&ms[v],e0,e1); ms[v]/=reps;
      CK(cudaGetLastError());
      long long pr[8]; cudaMemcpyFromSymbol(pr,g_prof,sizeof(pr)); cyc[v]=pr[1]-pr[0];
      res[v].resize(2*plane*np); CK(cudaMemcpy(res[v].data(),du,res[v].size()*4,cudaMemcpyDeviceToHost));
      cudaFree(du);
    }
    const int T=w+h-1+4;
    printf("np %4d %3dx%2d (%d steps):",np,w,h,T);
    const char* names[NV]={"lean","rows","v1","v1-nobar","v1-nocp","v1-nobar-nocp","warp"};
    for(int v=0;v<NV;++v){ size_t nd=0; for(size_t k=0;k<res[0].size();++k) if(res[v][k]!=res[0][k]) ++nd;
      printf("  %s %.1fus", names[v], ms[v]*1e3); if(v>=2&&v<6) printf(" (%.0f cyc/step)", (double)cyc[v]/T); printf(" [mis %zu]", nd); }
    printf("\n");
    cudaFree(dA);cudaFree(dB);
  }
  return 0;
}
