// TEST INFRASTRUCTURE (tests/test_finder_emulation.py): host-only check of build_finder() and of the algorithm behind
// wf_find_kernel + wf_replay_kernel (srt_wf_find.cuh): both are emulated on the built arrays in plain float and compared
// with an emulation of the shader's traversal (ray_intersects.glsl:99-133) in the same arithmetic -- rays from surface
// points, from outside, from a camera position, axis-parallel, with finite and infinite intersection_distance; and the
// OUTSIDE walk (distance bound theta, replay that reports rays it cannot be sure of).  The kernels themselves are checked
// bit for bit against the oracle by the GPU suite; this pins the argument they rest on where no GPU is needed.
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>
#include <time.h>
#include <algorithm>
#include <vector>
#include "srt_finder.h"
using namespace srt;
#include <string>
static float U2F(uint32_t u){float f;memcpy(&f,&u,4);return f;}
static uint32_t F2U(float f){uint32_t u;memcpy(&u,&f,4);return u;}
template<class T> std::vector<T> rd(const char* path){FILE* f=fopen(path,"rb");fseek(f,0,SEEK_END);long n=ftell(f);fseek(f,0,SEEK_SET);std::vector<T> v(n/sizeof(T));if(fread(v.data(),1,n,f)!=(size_t)n)abort();fclose(f);return v;}
static float ibox(const float*o,const float*inv,const float*mn,const float*mx){
  float tn=-INFINITY,tf=INFINITY; float t0[3],t1[3];
  for(int a=0;a<3;++a){t0[a]=(mn[a]-o[a])*inv[a];t1[a]=(mx[a]-o[a])*inv[a];}
  auto mn_=[](float x,float y){return (y<x)?y:x;}; auto mx_=[](float x,float y){return (x<y)?y:x;};
  tn=mx_(mx_(mn_(t0[0],t1[0]),mn_(t0[1],t1[1])),mn_(t0[2],t1[2]));
  tf=mn_(mn_(mx_(t0[0],t1[0]),mx_(t0[1],t1[1])),mx_(t0[2],t1[2]));
  return tn<=tf?((tn>=0)?tn:tf):INFINITY;
}
static bool tri(const TriRecord&r,const float*o,const float*d,float lim,float&t){
  const float e1[3]={r.e1.x,r.e1.y,r.e1.z},e2[3]={r.e2.x,r.e2.y,r.e2.z},v0[3]={r.v0.x,r.v0.y,r.v0.z};
  float h[3]={d[1]*e2[2]-d[2]*e2[1],d[2]*e2[0]-d[0]*e2[2],d[0]*e2[1]-d[1]*e2[0]};
  float a=e1[0]*h[0]+e1[1]*h[1]+e1[2]*h[2]; if(a>-0.0001f&&a<0.0001f)return false;
  float f=1.0f/a,s[3]={o[0]-v0[0],o[1]-v0[1],o[2]-v0[2]}; float u=f*(s[0]*h[0]+s[1]*h[1]+s[2]*h[2]); if(u<0||u>1)return false;
  float q[3]={s[1]*e1[2]-s[2]*e1[1],s[2]*e1[0]-s[0]*e1[2],s[0]*e1[1]-s[1]*e1[0]};
  float v=f*(d[0]*q[0]+d[1]*q[1]+d[2]*q[2]); if(v<0||u+v>1)return false;
  t=f*(e2[0]*q[0]+e2[1]*q[1]+e2[2]*q[2]); return t>0.00001f&&t<lim;
}
int main(int argc,char**argv){
  std::string dir=argv[1];
  auto bvhs=rd<srt_bvh>((dir+"/bvhs.bin").c_str()); auto nodes=rd<srt_bvh_node>((dir+"/nodes.bin").c_str());
  auto tris=rd<srt_triangle>((dir+"/tris.bin").c_str()); auto verts=rd<srt_vertex>((dir+"/verts.bin").c_str());
  uint32_t nt=tris.size();
  // ranks as relayout() assigns them
  std::vector<uint32_t> rank(nt,0); uint32_t nr=0;
  for(auto&b:bvhs){std::vector<uint32_t> st{b.first_index};while(!st.empty()){uint32_t i=st.back();st.pop_back();auto&n=nodes[i];if(n.prim_count>0){for(uint32_t k=0;k<n.prim_count;++k)rank[n.first_child_or_prim_index+k]=nr++;}else{st.push_back(n.first_child_or_prim_index);st.push_back(n.first_child_or_prim_index+1);}}}
  std::vector<TriRecord> tr(nt);
  for(uint32_t i=0;i<nt;++i){const float*v0=verts[tris[i].v0_idx].vertex,*v1=verts[tris[i].v1_idx].vertex,*v2=verts[tris[i].v2_idx].vertex;
    tr[i].v0=make_float4(v0[0],v0[1],v0[2],U2F(tris[i].material_idx));tr[i].e1=make_float4(v1[0]-v0[0],v1[1]-v0[1],v1[2]-v0[2],U2F(rank[i]));tr[i].e2=make_float4(v2[0]-v0[0],v2[1]-v0[1],v2[2]-v0[2],0);}
  FinderOut fo; const double t_b0=(double)clock()/CLOCKS_PER_SEC; struct timespec ts0,ts1; clock_gettime(CLOCK_MONOTONIC,&ts0); bool ok=build_finder(bvhs.data(),bvhs.size(),nodes.data(),nodes.size(),tris.data(),verts.data(),tr,fo);
  clock_gettime(CLOCK_MONOTONIC,&ts1); printf("build_finder: %.2f ms wall, %.2f ms cpu\n",(ts1.tv_sec-ts0.tv_sec)*1e3+(ts1.tv_nsec-ts0.tv_nsec)*1e-6,((double)clock()/CLOCKS_PER_SEC-t_b0)*1e3);
  printf("build ok=%d pairs=%zu ftris=%zu depth=%d cost ref %.1f finder %.1f worthwhile=%d\n",ok,fo.pairs.size(),fo.tris.size(),fo.depth,fo.cost_reference,fo.cost_finder,fo.worthwhile);
  if(!ok)return 0;
  std::vector<int> seen(nt,0); for(auto&r:fo.tris)seen[F2U(r.v0.w)]++; for(uint32_t i=0;i<nt;++i)if(seen[i]!=1){printf("triangle %u appears %d times\n",i,seen[i]);return 1;}
  // rays: from surface points and from outside
  uint64_t s=12345; auto rnd=[&](){s=s*6364136223846793005ull+1442695040888963407ull;return (float)((s>>33)&0xFFFFFF)/16777216.0f;};
  long mism=0,over=0,nray=200000; double cands=0; long n_out=0,o_mism=0,o_unc=0,o_over=0,o_cands=0,o_boxes=0,o_tris=0,o_rboxes=0;
  for(long it=0;it<nray;++it){
    uint32_t t0=(uint32_t)(rnd()*nt)%nt; float d[3]={2*rnd()-1,2*rnd()-1,2*rnd()-1};
    float a=rnd()*0.8f,b=rnd()*(0.9f-a); float o[3]={tr[t0].v0.x+a*tr[t0].e1.x+b*tr[t0].e2.x,tr[t0].v0.y+a*tr[t0].e1.y+b*tr[t0].e2.y,tr[t0].v0.z+a*tr[t0].e1.z+b*tr[t0].e2.z};
    if(it%3==1){for(int k=0;k<3;++k)o[k]-=40.0f*d[k];}
    if(it%4==2){  // from one camera position towards a point near the model (a frame's camera rays)
      auto&rb=nodes[bvhs[0].first_index]; float c[3],e[3]; for(int k=0;k<3;++k){c[k]=0.5f*(rb.min_bounds[k]+rb.max_bounds[k]);e[k]=rb.max_bounds[k]-rb.min_bounds[k];}
      float cam[3]={c[0],c[1],c[2]+2.2f*std::max(e[0],std::max(e[1],e[2]))};
      float tgt[3]={c[0]+(rnd()-0.5f)*1.3f*e[0],c[1]+(rnd()-0.5f)*1.3f*e[1],c[2]+(rnd()-0.5f)*e[2]};
      for(int k=0;k<3;++k){o[k]=cam[k];d[k]=tgt[k]-cam[k];}
      float nl=1.0f/sqrtf(d[0]*d[0]+d[1]*d[1]+d[2]*d[2]); for(int k=0;k<3;++k)d[k]*=nl*(0.5f+rnd()); }
    if(it%7==3)d[it%3]=0.0f;  // axis-parallel
    float tmax=(it%5==2)?rnd()*30.0f:INFINITY;
    float inv[3]={1.0f/d[0],1.0f/d[1],1.0f/d[2]};
    // reference
    float best=tmax; uint32_t hit=0xFFFFFFFFu;
    {std::vector<uint32_t> st{bvhs[0].first_index};while(!st.empty()){uint32_t n=st.back();st.pop_back();auto&nd=nodes[n];float bd=ibox(o,inv,nd.min_bounds,nd.max_bounds);if(bd<best&&!std::isinf(bd)){if(nd.prim_count>0){for(uint32_t k=0;k<nd.prim_count;++k){float t;if(tri(tr[nd.first_child_or_prim_index+k],o,d,best,t)){best=t;hit=nd.first_child_or_prim_index+k;}}}else{st.push_back(nd.first_child_or_prim_index);st.push_back(nd.first_child_or_prim_index+1);}}}}
    // finder
    struct C{uint32_t tri,rank;float t;}; std::vector<C> cs;
    {const float pad=fo.pad[0]; auto fb=[&](float mnx,float mny,float mnz,float mxx,float mxy,float mxz){float t0x=(mnx-o[0])*inv[0],t0y=(mny-o[1])*inv[1],t0z=(mnz-o[2])*inv[2],t1x=(mxx-o[0])*inv[0],t1y=(mxy-o[1])*inv[1],t1z=(mxz-o[2])*inv[2];float tn=fmaxf(fmaxf(fminf(t0x,t1x),fminf(t0y,t1y)),fminf(t0z,t1z)),tf=fminf(fminf(fmaxf(t0x,t1x),fmaxf(t0y,t1y)),fmaxf(t0z,t1z));return tn<=tf&&tf>=0&&tn<=tmax;};
      auto&rn=nodes[bvhs[0].first_index];
      if(fb(rn.min_bounds[0]-pad,rn.min_bounds[1]-pad,rn.min_bounds[2]-pad,rn.max_bounds[0]+pad,rn.max_bounds[1]+pad,rn.max_bounds[2]+pad)){
        std::vector<uint32_t> st{fo.root[0]};while(!st.empty()){uint32_t c=st.back();st.pop_back();if(c&SRT_LEAF_BIT){uint32_t i=c&~SRT_LEAF_BIT;for(;;){float t;if(tri(fo.tris[i],o,d,tmax,t))cs.push_back({F2U(fo.tris[i].v0.w),F2U(fo.tris[i].e1.w),t});if(!F2U(fo.tris[i].e2.w))break;++i;}}else{auto&p=fo.pairs[c];if(fb(p.a.x,p.a.y,p.a.z,p.a.w,p.b.x,p.b.y))st.push_back(p.d.z);if(fb(p.b.z,p.b.w,p.c.x,p.c.y,p.c.z,p.c.w))st.push_back(p.d.w);}}}}
    cands+=cs.size(); if(cs.size()>8)over++;
    std::sort(cs.begin(),cs.end(),[](const C&x,const C&y){return x.rank<y.rank;});
    float b2=tmax;uint32_t h2=0xFFFFFFFFu;std::vector<C> acc;
    for(auto&c:cs){if(!(c.t<b2))continue;bool sn=true;for(uint32_t n=fo.tri_leaf[c.tri];;){auto&R=fo.rnodes[n];float at=tmax;for(auto&a2:acc)if(a2.rank<F2U(R.b.w))at=a2.t;const float mn[3]={R.a.x,R.a.y,R.a.z},mx[3]={R.b.x,R.b.y,R.b.z};float bd=ibox(o,inv,mn,mx);if(!(bd<at&&!std::isinf(bd))){sn=false;break;}if(n==bvhs[0].first_index)break;n=F2U(R.a.w);if(n==0xFFFFFFFFu)break;}
      if(!sn)continue;b2=c.t;h2=c.tri;acc.push_back(c);}
    if(h2!=hit||(hit!=0xFFFFFFFFu&&b2!=best)){if(mism<5)printf("mismatch ray %ld: ref %u %g finder %u %g (cands %zu)\n",it,hit,best,h2,b2,cs.size());mism++;}
    // ---- OUTSIDE mode: origin outside the root box on some axis: nearest-first finder walk with distance culling
    {auto&rn=nodes[bvhs[0].first_index]; bool outside=false; for(int k=0;k<3;++k) if(o[k]<rn.min_bounds[k]||o[k]>rn.max_bounds[k]) outside=true;
     for(int k=0;k<3;++k) if(!(fabsf(inv[k])<INFINITY)) outside=false;  // axis-parallel rays are not handled as outside rays (srt_wf_find.cuh)
     if(outside){ n_out++;
      const float pad=fo.pad[0]; const float tpad=4.0f*pad/sqrtf(d[0]*d[0]+d[1]*d[1]+d[2]*d[2]);
      float tcull=tmax; std::vector<C> k2;
      auto fb=[&](float mnx,float mny,float mnz,float mxx,float mxy,float mxz,float&tn){float t0x=(mnx-o[0])*inv[0],t0y=(mny-o[1])*inv[1],t0z=(mnz-o[2])*inv[2],t1x=(mxx-o[0])*inv[0],t1y=(mxy-o[1])*inv[1],t1z=(mxz-o[2])*inv[2];tn=fmaxf(fmaxf(fminf(t0x,t1x),fminf(t0y,t1y)),fminf(t0z,t1z));float tf=fminf(fminf(fmaxf(t0x,t1x),fmaxf(t0y,t1y)),fmaxf(t0z,t1z));return tn<=tf&&tf>=0&&tn<=tcull;};
      float tn0;
      if(fb(rn.min_bounds[0]-pad,rn.min_bounds[1]-pad,rn.min_bounds[2]-pad,rn.max_bounds[0]+pad,rn.max_bounds[1]+pad,rn.max_bounds[2]+pad,tn0)){
        std::vector<uint32_t> st{fo.root[0]};while(!st.empty()){uint32_t c=st.back();st.pop_back();
          if(c&SRT_LEAF_BIT){uint32_t i=c&~SRT_LEAF_BIT;for(;;){float t;o_tris++;if(tri(fo.tris[i],o,d,tmax,t)){if(t<=tcull){k2.push_back({F2U(fo.tris[i].v0.w),F2U(fo.tris[i].e1.w),t});float nt=t+t*(1.0f/1024.0f)+tpad;if(nt<tcull)tcull=nt;}}if(!F2U(fo.tris[i].e2.w))break;++i;}}
          else{auto&p=fo.pairs[c];float tl,trr;o_boxes+=2;bool pl=fb(p.a.x,p.a.y,p.a.z,p.a.w,p.b.x,p.b.y,tl),pr=fb(p.b.z,p.b.w,p.c.x,p.c.y,p.c.z,p.c.w,trr);
            if(pl&&pr){if(tl<=trr){st.push_back(p.d.w);st.push_back(p.d.z);}else{st.push_back(p.d.z);st.push_back(p.d.w);}}else if(pl)st.push_back(p.d.z);else if(pr)st.push_back(p.d.w);}}}
      std::vector<C> kept; for(auto&c:k2) if(c.t<=tcull) kept.push_back(c);
      o_cands+=kept.size(); if(kept.size()>8){o_over++;}
      else{
      std::sort(kept.begin(),kept.end(),[](const C&x,const C&y){return x.rank<y.rank;});
      float b3=tmax;uint32_t h3=0xFFFFFFFFu;std::vector<C> acc3; bool uncertain=false;
      for(auto&c:kept){if(!(c.t<b3))continue;bool sn=true;for(uint32_t n=fo.tri_leaf[c.tri];;){auto&R=fo.rnodes[n];float at=tmax;for(auto&a2:acc3)if(a2.rank<F2U(R.b.w))at=a2.t;const float mn[3]={R.a.x,R.a.y,R.a.z},mx[3]={R.b.x,R.b.y,R.b.z};float bd=ibox(o,inv,mn,mx);o_rboxes++;if(!(bd<at&&!std::isinf(bd))){sn=false;break;}if(bd>=tcull)uncertain=true;if(n==bvhs[0].first_index)break;n=F2U(R.a.w);if(n==0xFFFFFFFFu)break;}
        if(!sn)continue;b3=c.t;h3=c.tri;acc3.push_back(c);}
      if(uncertain)o_unc++;
      else if(h3!=hit||(hit!=0xFFFFFFFFu&&b3!=best)){if(o_mism<5)printf("OUTSIDE mismatch ray %ld: ref %u %g finder %u %g (kept %zu)\n",it,hit,best,h3,b3,kept.size());o_mism++;}
      }}}
  }
  printf("outside rays %ld: mismatches %ld uncertain %ld overflow %ld kept/ray %.2f boxes/ray %.1f tris/ray %.1f replay boxes/ray %.1f\n",n_out,o_mism,o_unc,o_over,o_cands/(double)std::max(1L,n_out),o_boxes/(double)std::max(1L,n_out),o_tris/(double)std::max(1L,n_out),o_rboxes/(double)std::max(1L,n_out));
  printf("rays %ld mismatches %ld overflow(>8) %ld cands/ray %.2f\n",nray,mism,over,cands/nray);
  return mism!=0||o_mism!=0;
}
