// exactness harness (CPU, IEEE binary32 with fmaf): (1) magic-number floor of exp256, (2) division by a shared reciprocal
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
static inline float b2f(uint32_t u){float f;memcpy(&f,&u,4);return f;}
static inline uint32_t f2b(float f){uint32_t u;memcpy(&u,&f,4);return u;}
static volatile float MAGIC = 12582912.0f;
// ---- (1)
static uint64_t check_floor(uint32_t lo, uint32_t hi){ // over float bit patterns of x
  uint64_t bad=0; const float L2E=b2f(0x3fb8aa3bu);
  for(uint64_t u=lo;u<hi;u++){ float x=b2f((uint32_t)u); if(!(x==x)) continue;
    if(x>88.3762626647949f) x=88.3762626647949f; if(x<-88.3762626647949f) x=-88.3762626647949f;
    float fx=fmaf(x,L2E,0.5f);
    float fl=floorf(fx); int n=(int)fl;
    float tt=fx+MAGIC; float r=tt-MAGIC; int gt=r>fx; float fl2=gt? r-1.0f:r; int n2=(int)(f2b(tt))-0x4B400000-gt;
    if(f2b(fl)!=f2b(fl2)||n!=n2){ if(fl==0.0f&&fl2==0.0f&&n==n2) continue; bad++; if(bad<5) printf("floor mismatch x=%a fx=%a fl=%a fl2=%a n=%d n2=%d\n",x,fx,fl,fl2,n,n2);} }
  return bad; }
// ---- (2)
static inline float div3(float a,float s,float y){ float q0=a*y; float e0=fmaf(-q0,s,a); float q1=fmaf(e0,y,q0); float e1=fmaf(-q1,s,a); return fmaf(e1,y,q1);}
static inline float div2(float a,float s,float y){ float q0=a*y; float e0=fmaf(-q0,s,a); return fmaf(e0,y,q0);}
typedef struct{int id; uint64_t n; uint64_t bad3,bad2;}Job;
static uint64_t rng(uint64_t*s){ *s^=*s<<13; *s^=*s>>7; *s^=*s<<17; return *s; }
static void* run(void*p){ Job*j=p; uint64_t s=0x9E3779B97F4A7C15ull*(j->id+1); j->bad3=j->bad2=0;
  for(uint64_t i=0;i<j->n;i++){
    // sigma: exponent in [2^-10, 2^16], any mantissa; a = x - mu with x half-integer |x|<2^20, mu arbitrary nearby or far
    uint64_t r=rng(&s); uint32_t ms=(uint32_t)r&0x7fffff; int es=117+(int)((r>>23)%27); float sg=b2f(((uint32_t)es<<23)|ms);
    if((i&63)==0) sg=b2f(((uint32_t)es<<23)|0x7fffff);           // all-ones significands
    if((i&63)==1) sg=b2f(((uint32_t)es<<23));                    // powers of two
    r=rng(&s); float x=(float)((int)(r%2000001)-1000000)-0.5f; if((i&7)==0) x=(float)((int)(r%201)-100)-0.5f;
    r=rng(&s); float mu; switch(r&3){case 0: mu=b2f((uint32_t)(r>>8)&0xffffffffu); if(!(mu==mu)||fabsf(mu)>1e30f) mu=0.25f; break;
       case 1: mu=x+(float)((double)(r>>11)/9007199254740992.0-0.5)*8.0f*sg; break; case 2: mu=(float)((double)(r>>11)/9007199254740992.0-0.5)*60.0f; break; default: mu=x+b2f(0x33000000u+(uint32_t)((r>>8)%0x0c000000u)); }
    float a=x-mu; float y=1.0f/sg; float want=a/sg;
    if(f2b(div3(a,sg,y))!=f2b(want)){ if(!(want==want)) continue; j->bad3++; if(j->bad3<4) printf("div3 mismatch a=%a s=%a got=%a want=%a\n",a,sg,div3(a,sg,y),want);}
    if(f2b(div2(a,sg,y))!=f2b(want)) j->bad2++;
  } return 0;}
int main(int argc,char**argv){
  int T=8; pthread_t th[64]; 
  // (1) all floats: positive 0..0x7f800000, negative 0x80000000..0xff800000
  uint64_t bad=0; bad+=check_floor(0,0x7f800000u); bad+=check_floor(0x80000000u,0xff800000u);
  printf("magic floor: %llu mismatches over all finite binary32 x\n",(unsigned long long)bad);
  uint64_t N=argc>1?strtoull(argv[1],0,10):400000000ull; Job jobs[64];
  for(int i=0;i<T;i++){jobs[i].id=i;jobs[i].n=N/T;pthread_create(&th[i],0,run,&jobs[i]);}
  uint64_t b3=0,b2=0; for(int i=0;i<T;i++){pthread_join(th[i],0);b3+=jobs[i].bad3;b2+=jobs[i].bad2;}
  printf("division by shared reciprocal over %llu samples: 2 corrections: %llu mismatches; 1 correction: %llu mismatches\n",(unsigned long long)N,(unsigned long long)b3,(unsigned long long)b2);
  return 0;}
