// check: Markstein 2-FMA division by the constants the kernels use (24, 12: cell/half-cell indices;
// 144, 720, 1152: squared lengths of tile segments) == IEEE division, bit for bit (zeros compare equal)
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <stdint.h>
static inline uint64_t sm(uint64_t* s){uint64_t z=(*s+=0x9e3779b97f4a7c15ULL);z=(z^(z>>30))*0xbf58476d1ce4e5b9ULL;z=(z^(z>>27))*0x94d049bb133111ebULL;return z^(z>>31);}
static inline double divc(double x, double d, double c){ double q0=x*c; double r=fma(-d,q0,x); return fma(r,c,q0); }
int main(int argc,char**argv){ long n=atol(argv[1]); uint64_t s=atol(argv[2]); long bad=0;
 static const double D[5]={24.0,12.0,144.0,720.0,1152.0};
 volatile double dv[5]; double rc[5]; for(int j=0;j<5;j++){ dv[j]=D[j]; rc[j]=1.0/D[j]; }
 for(long i=0;i<n;i++){ uint64_t u=sm(&s); double x; int m=i%5;
   if(m==0) x=((double)(u>>11)/9007199254740992.0*1300.0)-100.0;
   else if(m==1){ // near multiples of 12
      long k=(long)(sm(&s)%60)-5; double base=(double)k*12.0; int e=(int)(sm(&s)%60); double eps=ldexp((double)((int64_t)(u>>40)-8388608)/8388608.0,-e); x=base+eps; }
   else if(m==2){ int e=(int)(sm(&s)%80)-40; x=ldexp(1.0+(double)(u>>12)/4503599627370496.0,e); if(u&1)x=-x; }
   else if(m==3){ // dot products (pos - end) . (px, py): near multiples of the divisor (u near 0, 1, k/divisor)
      int j=(int)(sm(&s)%5); long k=(long)(sm(&s)%5)-2; int e=(int)(sm(&s)%60); double eps=ldexp((double)((int64_t)(u>>40)-8388608)/8388608.0,-e); x=(double)k*D[j]+eps*D[j]; }
   else x=((double)(u>>11)/9007199254740992.0*2e6)-1e6;
   for(int j=0;j<5;j++){ double a=x/dv[j],b=divc(x,D[j],rc[j]); if(a!=b){ if(bad<5)printf("%g: x=%a %a %a\n",D[j],x,a,b); bad++; } }
 }
 printf("n=%ld bad=%ld\n",n,bad); return bad!=0; }
