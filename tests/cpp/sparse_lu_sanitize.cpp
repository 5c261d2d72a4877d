// Driver of tests/test_host_sanitizers.py: the host sparse LU and its orderings
// (rails_b200/host/SparseLU.cpp) under AddressSanitizer + UBSan -- the plan of
// SURVEY.md section 5 ("ASan/UBSan on the host shim").  Grids, a random matrix with
// zero diagonal entries and a dense row, the empty and the 1 x 1 matrix.
#include "SparseLU.hpp"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace RAILS;
static void grid(int g, int dim, std::vector<int>&rp, std::vector<int>&ci, std::vector<double>&v){
  int n=1; for(int d=0;d<dim;++d) n*=g; rp.assign(1,0);
  for(int r=0;r<n;++r){ int idx[3]; int t=r; for(int d=0;d<dim;++d){idx[d]=t%g;t/=g;}
    int stride=1; std::vector<std::pair<int,double>> row; row.push_back({r,-2.0*dim-0.5});
    for(int d=0;d<dim;++d){ if(idx[d]>0) row.push_back({r-stride,1.1}); if(idx[d]<g-1) row.push_back({r+stride,0.9}); stride*=g; }
    for(auto&e:row){ci.push_back(e.first);v.push_back(e.second);} rp.push_back((int)ci.size()); }
}
static int check(const char*name,int n,std::vector<int>&rp,std::vector<int>&ci,std::vector<double>&v,const int*order=nullptr){
  SparseLU lu; int info=lu.factor(n,rp.data(),ci.data(),v.data(),0.1,order);
  if(info){printf("%s: singular %d\n",name,info);return 1;}
  std::vector<double> b(n),x(n); for(int i=0;i<n;++i) x[i]=b[i]=std::sin(i*0.37)+1.1;
  for(int trans=0;trans<2;++trans){ std::vector<double> y=b; lu.solve(y.data(),1,n,trans!=0);
    double res=0,nb=0; std::vector<double> ax(n,0.0);
    for(int i=0;i<n;++i) for(int p=rp[i];p<rp[i+1];++p){ if(!trans) ax[i]+=v[p]*y[ci[p]]; else ax[ci[p]]+=v[p]*y[i]; }
    for(int i=0;i<n;++i){res+=(ax[i]-b[i])*(ax[i]-b[i]);nb+=b[i]*b[i];}
    printf("%s n=%d nnz=%lld trans=%d rel res %.2e\n",name,n,lu.nnz(),trans,std::sqrt(res/nb)); if(std::sqrt(res/nb)>1e-9) return 1; }
  return 0; }
int main(){ int bad=0; std::vector<int> rp,ci; std::vector<double> v;
  grid(40,2,rp,ci,v); bad|=check("grid40^2",1600,rp,ci,v);
  { auto o=nested_dissection_order(1600,rp.data(),ci.data(),50); bad|=check("grid40^2 nd",1600,rp,ci,v,o.data()); printf("fill md %lld\n",symbolic_fill(1600,rp.data(),ci.data(),minimum_degree_order(1600,rp.data(),ci.data()).data())); }
  rp.clear();ci.clear();v.clear(); grid(22,3,rp,ci,v); bad|=check("grid22^3",22*22*22,rp,ci,v);
  // random sparse with zero diagonal entries (pivoting) and a dense row
  rp.assign(1,0);ci.clear();v.clear(); int n=700; srand(3);
  for(int i=0;i<n;++i){ std::vector<int> cols; if(i%7) cols.push_back(i); for(int k=0;k<4;++k) cols.push_back(rand()%n); if(i==n-3) for(int j=0;j<n;j+=2) cols.push_back(j);
    cols.push_back((i+1)%n);
    for(int c:cols){ci.push_back(c);v.push_back((rand()%2000-1000)/500.0+ (c==i?5.0:0.0));} rp.push_back((int)ci.size()); }
  bad|=check("random700",n,rp,ci,v);
  // empty and 1x1
  { std::vector<int> r1={0,1},c1={0}; std::vector<double> v1={2.0}; bad|=check("1x1",1,r1,c1,v1); SparseLU e; int rp0[1]={0}; if(e.factor(0,rp0,nullptr,nullptr)) bad=1; }
  printf(bad?"FAILED\n":"OK\n"); return bad; }
