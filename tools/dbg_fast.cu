// debug harness: device vs host fast_score_px on a random image
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "../orb_slam2_test_b200/csrc/k_fast.cuh"
__global__ void k(const uint8_t* img, int w, int h, int* out){
  int x = blockIdx.x*blockDim.x+threadIdx.x, y = blockIdx.y;
  if (x<3||x>=w-3||y<3||y>=h-3) return;
  out[y*w+x] = fast_score_px(img+y*w+x, w, 7);
}
__global__ void k2(const uint8_t* img, int w, int h, int* out){
  __shared__ uint8_t win[66*68];
  // one block handles a 60x60 window at (bx*54, by*54)
  int x0 = blockIdx.x*54, y0 = blockIdx.y*54;
  for (int i=threadIdx.x;i<60*60;i+=blockDim.x){int y=i/60,x=i%60; int gx=min(x0+x,w-1), gy=min(y0+y,h-1); win[y*68+x]=img[gy*w+gx];}
  __syncthreads();
  for (int i=threadIdx.x;i<54*54;i+=blockDim.x){int y=i/54,x=i%54; int gx=x0+x+3, gy=y0+y+3; if(gx<w-3&&gy<h-3) out[gy*w+gx]=fast_score_px(win+(y+3)*68+x+3,68,7);}
}
int main(){
  int w=640,h=480; std::vector<uint8_t> img(w*h); srand(1);
  for(int y=0;y<h;y++)for(int x=0;x<w;x++){ int v = ((x/9+y/7)%2)*90 + 60 + rand()%12; img[y*w+x]=v; }
  uint8_t* d; int* o; cudaMalloc(&d,w*h); cudaMalloc(&o,w*h*4); cudaMemcpy(d,img.data(),w*h,cudaMemcpyHostToDevice);
  for(int mode=0;mode<2;mode++){
  cudaMemset(o,0,w*h*4);
  if(mode==0) k<<<dim3((w+127)/128,h),128>>>(d,w,h,o); else k2<<<dim3((w+53)/54,(h+53)/54),128>>>(d,w,h,o);
  std::vector<int> out(w*h); cudaMemcpy(out.data(),o,w*h*4,cudaMemcpyDeviceToHost);
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  int bad=0,n=0; for(int y=3;y<h-3;y++)for(int x=3;x<w-3;x++){ int r=fast_score_px(img.data()+y*w+x,w,7); n+=r>0; if(r!=out[y*w+x]){ if(bad<10) printf("mismatch (%d,%d) host %d dev %d\n",x,y,r,out[y*w+x]); bad++; } }
  printf("mode %d corners %d mismatches %d\n",mode,n,bad);}
}
