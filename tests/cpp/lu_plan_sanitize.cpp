// Driver of tests/test_host_sanitizers.py: the host-side planner of the device LU
// solves (plan_triangle in rails_b200/csrc/sptrsv.cu: levels, chains, shared-memory
// slots, windows) exercised through rails_lu_plan_emulate under AddressSanitizer.
// No CUDA call is made.
#include "SparseLU.hpp"
#include "rails_cuda.h"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace RAILS;
static void grid(int g, int dim, std::vector<int>&rp, std::vector<int>&ci, std::vector<double>&v){
  int n=1; for(int d=0;d<dim;++d) n*=g; rp.assign(1,0);
  for(int r=0;r<n;++r){ int idx[3]; int t=r; for(int d=0;d<dim;++d){idx[d]=t%g;t/=g;}
    int stride=1; ci.push_back(r); v.push_back(-2.0*dim-0.5);
    for(int d=0;d<dim;++d){ if(idx[d]>0){ci.push_back(r-stride);v.push_back(1.1);} if(idx[d]<g-1){ci.push_back(r+stride);v.push_back(0.9);} stride*=g; }
    rp.push_back((int)ci.size()); } }
static int run(const char*name,int n,std::vector<int>&rp,std::vector<int>&ci,std::vector<double>&v){
  SparseLU lu; if(lu.factor(n,rp.data(),ci.data(),v.data())) return 1;
  int bad=0;
  for(int trans=0;trans<2;++trans){
    std::vector<double> b(n*3),x; for(size_t i=0;i<b.size();++i) b[i]=std::sin(i*0.37)+1.1; x=b; std::vector<double> h=b;
    long long stats[12];
    if(rails_lu_plan_emulate(n,lu.Lp.data(),lu.Li.data(),lu.Lx.data(),lu.Up.data(),lu.Ui.data(),lu.Ux.data(),lu.pinv.data(),lu.q.data(),trans,x.data(),3,n,stats)) return 1;
    lu.solve(h.data(),3,n,trans!=0);
    double d=0,nn=0; for(size_t i=0;i<x.size();++i){d+=(x[i]-h[i])*(x[i]-h[i]);nn+=h[i]*h[i];}
    printf("%s n=%d trans=%d levels %lld/%lld chains %lld/%lld windows %lld/%lld diff %.2e\n",name,n,trans,stats[0],stats[6],stats[2],stats[8],stats[3],stats[9],std::sqrt(d/nn));
    if(std::sqrt(d/nn)>1e-10) bad=1; }
  return bad; }
int main(){ int bad=0; std::vector<int> rp,ci; std::vector<double> v;
  grid(70,2,rp,ci,v); bad|=run("grid70^2",4900,rp,ci,v);
  rp.clear();ci.clear();v.clear(); grid(20,3,rp,ci,v); bad|=run("grid20^3",8000,rp,ci,v);
  rp.clear();ci.clear();v.clear(); grid(110,2,rp,ci,v); bad|=run("grid110^2",12100,rp,ci,v);
  { std::vector<int> r1={0,1},c1={0}; std::vector<double> v1={2.0}; bad|=run("1x1",1,r1,c1,v1); }
  printf(bad?"FAILED\n":"OK\n"); return bad; }
