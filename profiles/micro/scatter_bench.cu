// micro-benchmark: what bounds the record scatter?  nvcc -O3 -arch=sm_100a scatter_bench.cu -o scatter_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define P 8192
#define CS 32
__device__ __forceinline__ uint64_t mix64(uint64_t x){x^=x>>30;x*=0xbf58476d1ce4e5b9ULL;x^=x>>27;x*=0x94d049bb133111ebULL;x^=x>>31;return x;}
__global__ void gen(ulonglong2*r,uint64_t R){uint64_t i=blockIdx.x*(uint64_t)blockDim.x+threadIdx.x;for(;i<R;i+=(uint64_t)gridDim.x*blockDim.x)r[i]=make_ulonglong2(mix64(i*2654435761ULL+12345),1);}
__global__ void hist(const ulonglong2*r,uint64_t R,uint32_t*h){uint64_t i=blockIdx.x*(uint64_t)blockDim.x+threadIdx.x;for(;i<R;i+=(uint64_t)gridDim.x*blockDim.x)atomicAdd(&h[mix64(r[i].x+0x9e3779b97f4a7c15ULL)>>51],1u);}
__global__ void scan(const uint32_t*h,uint64_t*o){uint64_t s=0;for(int i=0;i<P;i++){o[i]=s;s+=h[i];}}
// mode 0: atomic + store; 1: atomic only (pos saved coalesced); 2: store only (pos read coalesced); 3: atomic(non-returning) + nothing
template<int MODE,int U>
__global__ void __launch_bounds__(256) scatter(const ulonglong2*__restrict__ recs,uint64_t R,const uint64_t*__restrict__ offs,uint32_t*cur,ulonglong2*__restrict__ out,uint32_t*posbuf){
  uint64_t i=blockIdx.x*(uint64_t)blockDim.x+threadIdx.x;const uint64_t step=(uint64_t)gridDim.x*blockDim.x;
  for(;i<R;i+=step*U){
    ulonglong2 r[U];uint64_t pos[U];
#pragma unroll
    for(int u=0;u<U;u++){uint64_t j=i+u*step;r[u]=j<R?recs[j]:make_ulonglong2(~0ULL,0);if(MODE==2)pos[u]=j<R?posbuf[j]:0;}
#pragma unroll
    for(int u=0;u<U;u++){if(r[u].x==~0ULL)continue;uint32_t p=mix64(r[u].x+0x9e3779b97f4a7c15ULL)>>51;
      if(MODE==0||MODE==1)pos[u]=offs[p]+atomicAdd(&cur[p*CS],1u);
      if(MODE==3)atomicAdd(&cur[p*CS],1u);}
#pragma unroll
    for(int u=0;u<U;u++){if(r[u].x==~0ULL)continue;uint64_t j=i+u*step;
      if(MODE==0||MODE==2)out[pos[u]]=r[u];
      if(MODE==1)posbuf[j]=(uint32_t)pos[u];}
  }
}
// mode 4: tile-ranked two-level: CTA tile of T*U records, smem counts per partition, ONE global atomic per non-empty (tile,partition)
template<int U>
__global__ void __launch_bounds__(1024,1) scatter_tile(const ulonglong2*__restrict__ recs,uint64_t R,const uint64_t*__restrict__ offs,uint32_t*cur,ulonglong2*__restrict__ out){
  extern __shared__ uint32_t sm[]; uint32_t*cnt=sm; uint32_t*base=sm+P;
  const uint64_t tile=(uint64_t)1024*U;
  for(uint64_t t0=blockIdx.x*tile;t0<R;t0+=gridDim.x*tile){
    for(int p=threadIdx.x;p<P;p+=1024)cnt[p]=0; __syncthreads();
    ulonglong2 r[U];uint32_t rk[U];uint32_t pp[U];
#pragma unroll
    for(int u=0;u<U;u++){uint64_t j=t0+u*1024+threadIdx.x;r[u]=j<R?recs[j]:make_ulonglong2(~0ULL,0);}
#pragma unroll
    for(int u=0;u<U;u++){if(r[u].x==~0ULL)continue;pp[u]=mix64(r[u].x+0x9e3779b97f4a7c15ULL)>>51;rk[u]=atomicAdd(&cnt[pp[u]],1u);}
    __syncthreads();
    for(int p=threadIdx.x;p<P;p+=1024){uint32_t c=cnt[p];if(c)base[p]=atomicAdd(&cur[p*CS],c);}
    __syncthreads();
#pragma unroll
    for(int u=0;u<U;u++){if(r[u].x==~0ULL)continue;out[offs[pp[u]]+base[pp[u]]+rk[u]]=r[u];}
    __syncthreads();
  }
}

// pairing scatter: a record parks in a per-CTA shared-memory slot of its partition until a second record of the same
// partition arrives; the pair leaves as ONE 32-byte (full-sector) store at an even position taken with atomicAdd(cursor, 2).
// Leftover singles are written from the partition's tail at the end.  offs2[p] must be even; cnt[p] = records of p.
__device__ __forceinline__ void st32(ulonglong2*dst, ulonglong2 a, ulonglong2 b){
#if USE_V4
  asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};"::"l"(dst),"l"(a.x),"l"(a.y),"l"(b.x),"l"(b.y):"memory");
#else
  dst[0]=a;dst[1]=b;
#endif
}
template<int U>
__global__ void __launch_bounds__(1024,1) scatter_pair(const ulonglong2*__restrict__ recs,uint64_t R,const uint64_t*__restrict__ offs2,const uint32_t*__restrict__ cnt,uint32_t*cur,uint32_t*tail,ulonglong2*__restrict__ out){
  extern __shared__ __align__(16) unsigned char smraw[];
  ulonglong2*parked=(ulonglong2*)smraw; uint32_t*state=(uint32_t*)(smraw+(size_t)P*16);
  for(int p=threadIdx.x;p<P;p+=1024)state[p]=0; __syncthreads();
  uint64_t i=blockIdx.x*(uint64_t)1024+threadIdx.x;const uint64_t step=(uint64_t)gridDim.x*1024;
  for(;i<R;i+=step*U){
    ulonglong2 r[U];
#pragma unroll
    for(int u=0;u<U;u++){uint64_t j=i+u*step;r[u]=j<R?recs[j]:make_ulonglong2(~0ULL,0);}
#pragma unroll
    for(int u=0;u<U;u++){ if(r[u].x==~0ULL)continue; uint32_t p=mix64(r[u].x+0x9e3779b97f4a7c15ULL)>>51;
      for(;;){
        uint32_t st=atomicCAS(&state[p],0u,2u);
        if(st==0){parked[p]=r[u];__threadfence_block();atomicExch(&state[p],1u);break;}
        st=atomicCAS(&state[p],1u,2u);
        if(st==1){ulonglong2 o=parked[p];__threadfence_block();atomicExch(&state[p],0u);
          uint64_t pos=offs2[p]+atomicAdd(&cur[p*CS],2u); st32(out+pos,o,r[u]);break;}
      }
    }
  }
  __syncthreads();
  for(int p=threadIdx.x;p<P;p+=1024) if(state[p]==1){uint32_t k=atomicAdd(&tail[p],1u);out[offs2[p]+cnt[p]-1-k]=parked[p];}
}
__global__ void scan2(const uint32_t*h,uint64_t*o){uint64_t s=0;for(int i=0;i<P;i++){o[i]=s;s+=(h[i]+1)&~1u;}}
__global__ void check(const ulonglong2*out,const uint64_t*offs2,const uint32_t*cnt,unsigned long long*bad,unsigned long long*sum){
  int p=blockIdx.x; unsigned long long b=0,s=0; for(uint32_t k=threadIdx.x;k<cnt[p];k+=blockDim.x){ulonglong2 r=out[offs2[p]+k]; if((mix64(r.x+0x9e3779b97f4a7c15ULL)>>51)!=(uint64_t)p)b++; s+=r.x;}
  atomicAdd(bad,b);atomicAdd(sum,s);}
__global__ void sumin(const ulonglong2*r,uint64_t R,unsigned long long*sum){uint64_t i=blockIdx.x*(uint64_t)blockDim.x+threadIdx.x;unsigned long long s=0;for(;i<R;i+=(uint64_t)gridDim.x*blockDim.x)s+=r[i].x;atomicAdd(sum,s);}
int main(){
  const uint64_t R=66000000;ulonglong2*recs,*out;uint32_t*h,*cur,*posbuf;uint64_t*offs;
  cudaMalloc(&recs,R*16);cudaMalloc(&out,R*16+64);cudaMalloc(&h,P*4);cudaMalloc(&cur,P*CS*4);cudaMalloc(&offs,P*8);cudaMalloc(&posbuf,R*4);
  gen<<<148*16,256>>>(recs,R);cudaMemset(h,0,P*4);hist<<<148*16,256>>>(recs,R,h);scan<<<1,1>>>(h,offs);
  cudaEvent_t e0,e1;cudaEventCreate(&e0);cudaEventCreate(&e1);float ms;
  cudaFuncSetAttribute(scatter_tile<8>,cudaFuncAttributeMaxDynamicSharedMemorySize,P*8);
  cudaFuncSetAttribute(scatter_tile<16>,cudaFuncAttributeMaxDynamicSharedMemorySize,P*8);
#define RUN(name,launch) for(int it=0;it<3;it++){cudaMemset(cur,0,P*CS*4);cudaEventRecord(e0);launch;cudaEventRecord(e1);cudaEventSynchronize(e1);cudaEventElapsedTime(&ms,e0,e1);if(it==2)printf("%-34s %.3f ms  (%s)\n",name,ms,cudaGetErrorString(cudaGetLastError()));}
  RUN("atomic+store U4 g2368",(scatter<0,4><<<148*16,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("atomic only (save pos)",(scatter<1,4><<<148*16,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("store only (read pos)",(scatter<2,4><<<148*16,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("RED only (non-returning)",(scatter<3,4><<<148*16,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("atomic+store U8",(scatter<0,8><<<148*8,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("atomic+store U2 g4736",(scatter<0,2><<<148*32,256>>>(recs,R,offs,cur,out,posbuf)))
  RUN("tile-ranked 8192/tile",(scatter_tile<8><<<148,1024,P*8>>>(recs,R,offs,cur,out)))
  RUN("tile-ranked 16384/tile",(scatter_tile<16><<<148,1024,P*8>>>(recs,R,offs,cur,out)))
  { uint64_t*offs2;uint32_t*tail;unsigned long long*chk;cudaMalloc(&offs2,P*8);cudaMalloc(&tail,P*4);cudaMalloc(&chk,24);cudaFree(out);cudaMalloc(&out,(R+P+8)*16);
    scan2<<<1,1>>>(h,offs2); size_t sm=(size_t)P*16+P*4;
    cudaFuncSetAttribute(scatter_pair<4>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm);cudaFuncSetAttribute(scatter_pair<2>,cudaFuncAttributeMaxDynamicSharedMemorySize,(int)sm);
    for(int it=0;it<3;it++){cudaMemset(cur,0,P*CS*4);cudaMemset(tail,0,P*4);cudaEventRecord(e0);scatter_pair<4><<<148,1024,sm>>>(recs,R,offs2,h,cur,tail,out);cudaEventRecord(e1);cudaEventSynchronize(e1);cudaEventElapsedTime(&ms,e0,e1);if(it==2)printf("%-34s %.3f ms  (%s)\n","pairing scatter U4 (32 B stores)",ms,cudaGetErrorString(cudaGetLastError()));}
    cudaMemset(chk,0,24);check<<<P,128>>>(out,offs2,h,chk,chk+1);sumin<<<148*8,256>>>(recs,R,chk+2);unsigned long long hc[3];cudaMemcpy(hc,chk,24,cudaMemcpyDeviceToHost);printf("pairing check: misplaced=%llu checksum %s\n",hc[0],hc[1]==hc[2]?"ok":"MISMATCH");
    for(int it=0;it<3;it++){cudaMemset(cur,0,P*CS*4);cudaMemset(tail,0,P*4);cudaEventRecord(e0);scatter_pair<2><<<148,1024,sm>>>(recs,R,offs2,h,cur,tail,out);cudaEventRecord(e1);cudaEventSynchronize(e1);cudaEventElapsedTime(&ms,e0,e1);if(it==2)printf("%-34s %.3f ms  (%s)\n","pairing scatter U2",ms,cudaGetErrorString(cudaGetLastError()));}
  }
  // plain copy for reference
  RUN("copy 1.06 GB (memcpy d2d)",(cudaMemcpyAsync(out,recs,R*16,cudaMemcpyDeviceToDevice)))
  return 0;
}
