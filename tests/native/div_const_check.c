// check: Markstein 2-FMA division by 24 and 12 == IEEE division, then floor
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <stdint.h>
static inline uint64_t sm(uint64_t* s){uint64_t z=(*s+=0x9e3779b97f4a7c15ULL);z=(z^(z>>30))*0xbf58476d1ce4e5b9ULL;z=(z^(z>>27))*0x94d049bb133111ebULL;return z^(z>>31);}
static inline double divc(double x, double d, double c){ double q0=x*c; double r=fma(-d,q0,x); return fma(r,c,q0); }
int main(int argc,char**argv){ long n=atol(argv[1]); uint64_t s=atol(argv[2]); long bad=0;
 volatile double d24=24.0,d12=12.0; double c24=1.0/24.0,c12=1.0/12.0;
 for(long i=0;i<n;i++){ uint64_t u=sm(&s); double x; int m=i%4;
   if(m==0) x=((double)(u>>11)/9007199254740992.0*1300.0)-100.0;
   else if(m==1){ // near multiples of 24/12
      long k=(long)(sm(&s)%60)-5; double base=(double)k*12.0; int e=(int)(sm(&s)%60); double eps=ldexp((double)((int64_t)(u>>40)-8388608)/8388608.0,-e); x=base+eps; }
   else if(m==2){ int e=(int)(sm(&s)%80)-40; x=ldexp(1.0+(double)(u>>12)/4503599627370496.0,e); if(u&1)x=-x; }
   else x=((double)(u>>11)/9007199254740992.0*2e6)-1e6;
   double a=x/d24,b=divc(x,24.0,c24); if(a!=b){ if(bad<5)printf("24: x=%a %a %a\n",x,a,b); bad++; }
   a=x/d12; b=divc(x,12.0,c12); if(a!=b){ if(bad<5)printf("12: x=%a %a %a\n",x,a,b); bad++; }
 }
 printf("n=%ld bad=%ld\n",n,bad); return bad!=0; }
