#include <immintrin.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <cfenv>
#include <unistd.h>
#include <sys/syscall.h>
struct tilecfg { uint8_t palette, start_row; uint8_t rsv[14]; uint16_t colsb[16]; uint8_t rows[16]; };
static float bf2f(uint16_t b){ uint32_t u=(uint32_t)b<<16; float f; memcpy(&f,&u,4); return f;}
static uint16_t f2bf(float f){ uint32_t u; memcpy(&u,&f,4); u += 0x7fff + ((u>>16)&1); return u>>16; }
static float rnd(){ return ((rand()%2001)-1000)/ (float)(1+rand()%977); }
int main(){
  if (syscall(SYS_arch_prctl, 0x1023, 18)) { printf("denied\n"); return 1; }
  tilecfg cfg; memset(&cfg,0,sizeof cfg); cfg.palette=1;
  for(int i=0;i<8;i++){ cfg.colsb[i]=64; cfg.rows[i]=16; }
  _tile_loadconfig(&cfg);
  alignas(64) uint16_t A[16][32], B[16][32]; alignas(64) float C[16][16];
  srand(2);
  for (int npairs=1; npairs<=3; npairs++){
    int n=0, mRN=0, mRZ=0, mseqpairRN=0; int shown=0;
    for (int trial=0; trial<300; trial++){
      memset(A,0,sizeof A); memset(B,0,sizeof B);
      for(int i=0;i<16;i++)for(int j=0;j<2*npairs;j++) A[i][j]=f2bf(rnd());
      for(int k=0;k<npairs;k++)for(int j=0;j<32;j++) B[k][j]=f2bf(rnd());
      for(int i=0;i<16;i++)for(int j=0;j<16;j++) C[i][j]= rnd();
      float C0[16][16]; memcpy(C0,C,sizeof C);
      _tile_loadd(0, C, 64); _tile_loadd(1, A, 64); _tile_loadd(2, B, 64);
      _tile_dpbf16ps(0,1,2);
      _tile_stored(0, C, 64);
      for(int m=0;m<16;m++)for(int nn=0;nn<16;nn++){
        __float128 e=C0[m][nn];
        for(int k=0;k<npairs;k++){ e += (__float128)bf2f(A[m][2*k])*bf2f(B[k][2*nn]); e += (__float128)bf2f(A[m][2*k+1])*bf2f(B[k][2*nn+1]); }
        fesetround(FE_TONEAREST); float rn=(float)e;
        // RZ
        float rz = rn; if (fabsl((long double)rz) > fabsl((long double)e)) rz = nextafterf(rz, 0.0f);
        float q=C0[m][nn];
        for(int k=0;k<npairs;k++){ __float128 t = (__float128)bf2f(A[m][2*k+1])*bf2f(B[k][2*nn+1]) + (__float128)bf2f(A[m][2*k])*bf2f(B[k][2*nn]); q = (float)((__float128)q + t); }
        n++; if(rn!=C[m][nn]) mRN++; if(rz!=C[m][nn]) mRZ++; if(q!=C[m][nn]) mseqpairRN++;
        if (npairs==1 && rn!=C[m][nn] && shown<5){ shown++; printf("  c0=%a a0=%a b0=%a a1=%a b1=%a  hw=%a rn=%a\n", C0[m][nn], bf2f(A[m][0]),bf2f(B[0][2*nn]),bf2f(A[m][1]),bf2f(B[0][2*nn+1]), C[m][nn], rn); }
      }
    }
    printf("npairs=%d n=%d mism exactRN=%d exactRZ=%d seqpairRN=%d\n",npairs,n,mRN,mRZ,mseqpairRN);
  }
  _tile_release();
}
