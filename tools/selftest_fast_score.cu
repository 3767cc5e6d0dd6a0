
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <algorithm>
#include <cuda_runtime.h>
#ifndef __CUDA_ARCH__
using std::min; using std::max;
#endif
__host__ __device__ __forceinline__ int fast_score_px(const uint8_t* __restrict__ p, int stride, int t) {
    const int c = p[0];
    const int hi = c + t, lo = c - t;
    // cardinal early-out (KFAST.h:186-235): any 9-arc contains two adjacent compass points
    const int p9 = p[3 * stride], p5 = p[3], p1 = p[-3 * stride], p13 = p[-3];
    const bool b = ((p9 > hi) & (p5 > hi)) | ((p5 > hi) & (p1 > hi)) | ((p1 > hi) & (p13 > hi)) | ((p13 > hi) & (p9 > hi));
    const bool d = ((p9 < lo) & (p5 < lo)) | ((p5 < lo) & (p1 < lo)) | ((p1 < lo) & (p13 < lo)) | ((p13 < lo) & (p9 < lo));
    if (!(b | d)) return 0;
    int ring[16];
    ring[0] = p9;                 ring[1] = p[3 * stride + 1];  ring[2] = p[2 * stride + 2];  ring[3] = p[stride + 3];
    ring[4] = p5;                 ring[5] = p[-stride + 3];     ring[6] = p[-2 * stride + 2]; ring[7] = p[-3 * stride + 1];
    ring[8] = p1;                 ring[9] = p[-3 * stride - 1]; ring[10] = p[-2 * stride - 2]; ring[11] = p[-stride - 3];
    ring[12] = p13;               ring[13] = p[stride - 3];     ring[14] = p[2 * stride - 2]; ring[15] = p[3 * stride - 1];
    uint32_t mb = 0, md = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) { mb |= (uint32_t)(ring[k] > hi) << k; md |= (uint32_t)(ring[k] < lo) << k; }
    // 9 contiguous set bits in the circular 16-bit mask
    auto run9 = [](uint32_t m) {
        m |= m << 16;
        uint32_t a = m & (m >> 1);
        a &= a >> 2;
        a &= a >> 4;
        a &= m >> 8;
        return (a & 0xFFFFu) != 0;
    };
    if (!(run9(mb) | run9(md))) return 0;
    // score = max over the 16 arcs of 9 of max(min_k(c - x_k), min_k(x_k - c))
    //       = max(c - min_arcs(max_k x_k), max_arcs(min_k x_k) - c).
    // Sliding max/min of width 9 on the raw ring values (all in [0,255], non-negative: ptxas 12.9
    // mis-packs *signed* 32-bit min/max chains into VIMNMX3.U16x2 at -O1+):
    // width 8 by three doubling steps, then one more element.
    int mn1[16], mx1[16], mn2[16], mx2[16], mn4[16], mx4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn1[k] = min(ring[k], ring[(k + 1) & 15]); mx1[k] = max(ring[k], ring[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn2[k] = min(mn1[k], mn1[(k + 2) & 15]); mx2[k] = max(mx1[k], mx1[(k + 2) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; ++k) { mn4[k] = min(mn2[k], mn2[(k + 4) & 15]); mx4[k] = max(mx2[k], mx2[(k + 4) & 15]); }
    int lo_of_max = 255, hi_of_min = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo_of_max = min(lo_of_max, max(mx4[k], ring[(k + 8) & 15]));
        hi_of_min = max(hi_of_min, min(mn4[k], ring[(k + 8) & 15]));
    }
    const int dark = c - lo_of_max, bright = hi_of_min - c;
    const int best = dark > bright ? dark : bright;
    return best & 0xFF;
}


__global__ void k(const uint8_t* img, int w, int h, int t, uint8_t* out){
  int x = blockIdx.x*blockDim.x+threadIdx.x, y = blockIdx.y;
  if (x>=3 && x<w-3 && y>=3 && y<h-3) out[y*w+x] = fast_score_px(img+y*w+x, w, t); }
__global__ void ks(const uint8_t* img, int w, int h, int t, uint8_t* out){   // via shared memory, stride 80
  __shared__ uint8_t px[24*80];
  int x0 = blockIdx.x*64, y0 = blockIdx.y*16;
  for (int i=threadIdx.x;i<24*72;i+=blockDim.x){int r=i/72,c=i%72; int gy=min(max(y0-4+r,0),h-1), gx=min(max(x0-4+c,0),w-1); px[r*80+c]=img[gy*w+gx];}
  __syncthreads();
  for (int i=threadIdx.x;i<16*64;i+=blockDim.x){int r=i/64,c=i%64; int gx=x0+c, gy=y0+r;
    if (gx>=3 && gx<w-3 && gy>=3 && gy<h-3) out[gy*w+gx]=fast_score_px(&px[(r+4)*80+c+4],80,t);} }
int main(){ int w=640,h=480; uint8_t* img=(uint8_t*)malloc(w*h); srand(1);
 for(int i=0;i<w*h;i++) img[i]=(rand()%4==0)?rand()%256:((i/w/13+i%w/17)%2?200:60);
 uint8_t *d,*o,*o2; cudaMalloc(&d,w*h); cudaMalloc(&o,w*h); cudaMalloc(&o2,w*h); cudaMemset(o,0,w*h); cudaMemset(o2,0,w*h); cudaMemcpy(d,img,w*h,cudaMemcpyHostToDevice);
 k<<<dim3((w+127)/128,h),128>>>(d,w,h,40,o); ks<<<dim3(w/64,h/16),256>>>(d,w,h,40,o2);
 uint8_t* g=(uint8_t*)malloc(w*h); uint8_t* g2=(uint8_t*)malloc(w*h); cudaMemcpy(g,o,w*h,cudaMemcpyDeviceToHost); cudaMemcpy(g2,o2,w*h,cudaMemcpyDeviceToHost);
 printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
 int bad=0,bad2=0,n=0; for(int y=3;y<h-3;y++)for(int x=3;x<w-3;x++){int s=fast_score_px(img+y*w+x,w,40); if(s)n++; if(s!=g[y*w+x]){if(bad<5)printf("glob %d %d: host %d dev %d\n",x,y,s,g[y*w+x]);bad++;} if(s!=g2[y*w+x]){if(bad2<5)printf("smem %d %d: host %d dev %d\n",x,y,s,g2[y*w+x]);bad2++;}}
 printf("corners %d bad_global %d bad_smem %d\n",n,bad,bad2); return 0;}
