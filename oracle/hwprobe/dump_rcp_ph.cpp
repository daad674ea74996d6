#include <immintrin.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
int main(){
  static uint16_t out[65536];
  for (int base=0; base<65536; base+=32){
    uint16_t in[32]; for(int i=0;i<32;i++) in[i]=base+i;
    __m512h x = _mm512_loadu_ph(in);
    __m512h r = _mm512_rcp_ph(x);
    _mm512_storeu_ph(out+base, r);
  }
  // binade invariance for normal inputs exp 1..30 whose result is normal
  int bad=0, total=0;
  for (int e=1;e<=30;e++) for(int m=0;m<1024;m++){
    uint16_t x=(e<<10)|m; uint16_t r=out[x];
    uint16_t x0=(15<<10)|m; uint16_t r0=out[x0]; // 1.m
    int re=(r>>10)&31, r0e=(r0>>10)&31;
    if (re==0||re==31) continue;
    total++;
    // expected: exponent shifts by (15-e)
    int expe = r0e + (15-e);
    if ((r&1023)!=(r0&1023) || re!=expe) bad++;
  }
  printf("binade-variant entries: %d of %d\n", bad,total);
  // compare with RN(1/x)
  int diff=0, n=0;
  for (int x=0x5fbc; x<=0x7260; x++){ // ~495..13000
    _Float16 h; uint16_t xx=x; memcpy(&h,&xx,2);
    _Float16 q = (_Float16)(1.0f/(float)h); uint16_t qb; memcpy(&qb,&q,2);
    n++; if(qb!=out[x]) diff++;
  }
  printf("range diff vs RN: %d of %d\n", diff,n);
  FILE*f=fopen("rcp_ph_mant.bin","wb"); fwrite(out+(15<<10),2,1024,f); fclose(f);
  f=fopen("rcp_ph_all.bin","wb"); fwrite(out,2,65536,f); fclose(f);
  return 0;
}
