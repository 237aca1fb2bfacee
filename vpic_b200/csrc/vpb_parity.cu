// vpb_parity.cu -- PARITY deposit mode: every float / double sum of the step is
// formed in the reference's own order, so accumulators, currents, fields and
// energy diagnostics are BIT-IDENTICAL to the reference's scalar build run with
// --tpp P (P = aa->n_pipeline for advance_p, vpic_b200_set_reference_pipelines
// for the diagnostics).
//
// Reference order of advance_p (pipeline/advance_p_pipeline.cc:54-82,245-314):
// pipeline r pushes particles DISTRIBUTE(np,16,r,P) in array order into its
// private accumulator copy r+1, the host pushes the np%16 stragglers into copy
// 0; every deposit is "a[voxel].j += v", streak after streak; species follow
// each other on the same copies; reduce_accumulator_array then adds the copies
// with a fixed tree (reduce_accumulators_pipeline.cc:86-102).
//
// Device algorithm: push every particle twice.  Pass 1 counts each particle's
// streaks (1 if it stays in its voxel); an exclusive scan gives every streak a
// record slot, so record index order == (particle, streak) order == the
// reference's deposit order.  Pass 2 writes the particles and one record
// (destination bin = copy*nv + voxel, 12 currents) per streak.  Records are
// then grouped by bin with the counting-sort front end, each bin's record
// indices are put in ascending order by one warp, and 12 lanes add the records
// to the accumulator one after the other.  Roughly 3x the traffic and 2x the
// arithmetic of the FAST kernel: a verification mode, not the fast path.
#include "vpb_push.cuh"
#include "vpb_sortutil.cuh"
#include "vpb_hydro.cuh"

namespace vpb {

constexpr int PAR_MAXP = 15;     // pipelines whose order can be reproduced

struct ParRanges { int P; int start[PAR_MAXP+2]; };   // pipeline r owns [start[r], start[r+1]); host owns [start[P], np)

__device__ __forceinline__ int copy_of( const ParRanges & R, int n ) {
  if( n>=R.start[R.P] ) return 0;
  int r = 0;
  while( n>=R.start[r+1] ) r++;
  return r+1;
}

template<bool COUNT_ONLY>
__global__ void __launch_bounds__(128)
k_parity_push( vpb_particle_t * __restrict__ p, int np, const float * __restrict__ ip,
               int * __restrict__ cnt, const int * __restrict__ off,
               int * __restrict__ rec_bin, float * __restrict__ rec_val,
               vpb_particle_mover_t * __restrict__ pm, int max_nm, int * __restrict__ counters,
               GridDev g, PushCoef c, ParRanges R ) {
  const int n = blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  float4 * pp = reinterpret_cast<float4 *>( p + n );
  const float4 r0 = pp[0], r1 = pp[1];
  const float dx = r0.x, dy = r0.y, dz = r0.z; const int ii = __float_as_int( r0.w );
  float ux = r1.x, uy = r1.y, uz = r1.z;
  const Interp f = load_interp( ip, ii );
  boris( f, c, dx, dy, dz, ux, uy, uz );
  const float sux = ux, suy = uy, suz = uz;
  const float one = 1.0f, one_third = 1.0/3.0;
  float v0 = one/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
  ux *= c.cdt_dx; uy *= c.cdt_dy; uz *= c.cdt_dz;
  ux *= v0; uy *= v0; uz *= v0;
  v0 = dx + ux; const float v1 = dy + uy, v2 = dz + uz;
  const float v3 = v0 + ux, v4 = v1 + uy, v5 = v2 + uz;
  const float q = r1.w*c.qsp;
  const int bin0 = COUNT_ONLY ? 0 : copy_of( R, n )*g.nv;
  float j[12];
  if( v3<=one && v4<=one && v5<=one && -v3<=one && -v4<=one && -v5<=one ) {
    if( COUNT_ONLY ) { cnt[n] = 1; return; }
    pp[0] = make_float4( v3, v4, v5, r0.w );
    pp[1] = make_float4( sux, suy, suz, r1.w );
    streak_currents( j, q, v0, v1, v2, ux, uy, uz, q*ux*uy*uz*one_third );
    const int s = off[n];
    rec_bin[s] = bin0 + ii;
    float4 * out = reinterpret_cast<float4 *>( rec_val + 12*(size_t)s );
    out[0] = make_float4( j[0], j[1], j[2], j[3] ); out[1] = make_float4( j[4], j[5], j[6], j[7] ); out[2] = make_float4( j[8], j[9], j[10], j[11] );
    return;
  }
  MoveState ms;
  ms.dx = dx; ms.dy = dy; ms.dz = dz; ms.i = ii;
  ms.ux = sux; ms.uy = suy; ms.uz = suz;
  ms.dispx = ux; ms.dispy = uy; ms.dispz = uz;
  int status, k = 0;
  const int s0 = COUNT_ONLY ? 0 : off[n];
  do {
    int vox;
    status = move_p_streak( ms, q, g, j, vox );
    if( !COUNT_ONLY ) {
      rec_bin[s0+k] = bin0 + vox;
      float4 * out = reinterpret_cast<float4 *>( rec_val + 12*(size_t)( s0+k ) );
      out[0] = make_float4( j[0], j[1], j[2], j[3] ); out[1] = make_float4( j[4], j[5], j[6], j[7] ); out[2] = make_float4( j[8], j[9], j[10], j[11] );
    }
    k++;
  } while( status==1 );
  if( COUNT_ONLY ) { cnt[n] = k; return; }
  pp[0] = make_float4( ms.dx, ms.dy, ms.dz, __int_as_float( ms.i ) );
  pp[1] = make_float4( ms.ux, ms.uy, ms.uz, r1.w );
  if( status==2 ) {
    const int slot = atomicAdd( counters+0, 1 );
    if( slot<max_nm ) reinterpret_cast<float4 *>( pm )[slot] = make_float4( ms.dispx, ms.dispy, ms.dispz, __int_as_float( n ) );
    else atomicAdd( counters+1, 1 );
  }
}

// One warp per bin: order the bin's record indices ascending, then lanes 0..11
// add the records to the accumulator sequentially ("a[voxel].j += v").
__global__ void __launch_bounds__(32*FIX_WARPS)
k_parity_sum( const int * __restrict__ start, int nbin, int nrec, int * __restrict__ perm, int * __restrict__ big_scratch,
              const float * __restrict__ rec_val, float * __restrict__ acc0, float * __restrict__ acc_priv, int nv ) {
  __shared__ int s[FIX_WARPS][FIX_SMEM];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int bin = blockIdx.x*FIX_WARPS + w;
  if( bin>=nbin ) return;
  const int b = start[bin], e = bin+1<nbin ? start[bin+1] : nrec, n = e-b;
  if( n<=0 ) return;
  int n2 = 32; while( n2<n ) n2 <<= 1;
  int * a = n2<=FIX_SMEM ? s[w] : big_scratch + 2*(size_t)b;
  for( int t=lane; t<n2; t+=32 ) a[t] = t<n ? perm[b+t] : 0x7fffffff;
  __threadfence_block(); __syncwarp();
  bitonic_mem( a, n2, lane );
  if( lane<12 ) {
    const int copy = bin/nv, voxel = bin - copy*nv;
    float * dst = ( copy==0 ? acc0 : acc_priv + (size_t)( copy-1 )*nv*ACC_STRIDE ) + (size_t)voxel*ACC_STRIDE + lane;
    float v = *dst;
    for( int t=0; t<n; t++ ) v = v + rec_val[ 12*(size_t)a[t] + lane ];
    *dst = v;
  }
}

// reduce_accumulators_pipeline.cc:86-102
__global__ void __launch_bounds__(256)
k_parity_reduce( float * __restrict__ a0, const float * __restrict__ priv, int nr, size_t n ) {
  const size_t k = (size_t)blockIdx.x*blockDim.x + threadIdx.x;
  if( k>=n ) return;
  float v = a0[k];
# define B(r) priv[ (size_t)(r)*n + k ]
  switch( nr ) {
  case 0: break;
  case 1: v =     v+B(0); break;
  case 2: v =    (v+B(0))+ B(1); break;
  case 3: v =    (v+B(0))+(B(1)+B(2)); break;
  case 4: v =   ((v+B(0))+(B(1)+B(2)))+ B(3); break;
  case 5: v =   ((v+B(0))+(B(1)+B(2)))+(B(3)+B(4)); break;
  case 6: v =   ((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+ B(5)); break;
  case 7: v =   ((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))); break;
  case 8: v =  (((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))))+ B(7); break;
  default: v = (((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))))+(B(7)+B(8)); break;
  }
# undef B
  a0[k] = v;
}

// ---- diagnostics in the reference's summation order -------------------------
// energy_f_pipeline.c:13-95 / energy_p_pipeline.cc:17-115: each pipeline adds
// its share (blocks of 16 items, DISTRIBUTE) into a double in item order, the
// host adds the remainder, the partial sums are added in pipeline order.

template<bool VACUUM>
__global__ void __launch_bounds__(256)
k_energy_f_values( double * __restrict__ val, const float * __restrict__ f, const int16_t * __restrict__ mat,
                   const vpb_material_coefficient_t * __restrict__ m, GridDev g ) {
  const int ncell = g.nx*g.ny*g.nz;
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if( i>=ncell ) return;
  const size_t P = (size_t)g.nvp;
  const float * ex = f + F_EX*P, * ey = f + F_EY*P, * ez = f + F_EZ*P;
  const float * bx = f + F_CBX*P, * by = f + F_CBY*P, * bz = f + F_CBZ*P;
  const int x = i % g.nx, t = i / g.nx, y = t % g.ny, z = t / g.ny;
  const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const int vx = v+1, vy = v+g.sy, vz = v+g.sz, vyz = v+g.sy+g.sz, vzx = v+g.sz+1, vxy = v+1+g.sy;
  double e[6];
  if( VACUUM ) {
    const float qepsx = 0.25*m->epsx, qepsy = 0.25*m->epsy, qepsz = 0.25*m->epsz;
    const float hrmux = 0.50*m->rmux, hrmuy = 0.50*m->rmuy, hrmuz = 0.50*m->rmuz;
    e[0] = (double)( qepsx*( ex[v]*ex[v] + ex[vy]*ex[vy] + ex[vz]*ex[vz] + ex[vyz]*ex[vyz] ) );
    e[1] = (double)( qepsy*( ey[v]*ey[v] + ey[vz]*ey[vz] + ey[vx]*ey[vx] + ey[vzx]*ey[vzx] ) );
    e[2] = (double)( qepsz*( ez[v]*ez[v] + ez[vx]*ez[vx] + ez[vy]*ez[vy] + ez[vxy]*ez[vxy] ) );
    e[3] = (double)( hrmux*( bx[v]*bx[v] + bx[vx]*bx[vx] ) );
    e[4] = (double)( hrmuy*( by[v]*by[v] + by[vy]*by[vy] ) );
    e[5] = (double)( hrmuz*( bz[v]*bz[v] + bz[vz]*bz[vz] ) );
  } else {
    const int16_t * emx = mat + 0*P, * emy = mat + 1*P, * emz = mat + 2*P;
    const int16_t * fmx = mat + 4*P, * fmy = mat + 5*P, * fmz = mat + 6*P;
    e[0] = 0.25*(double)( m[emx[v]].epsx*ex[v]*ex[v] + m[emx[vy]].epsx*ex[vy]*ex[vy] + m[emx[vz]].epsx*ex[vz]*ex[vz] + m[emx[vyz]].epsx*ex[vyz]*ex[vyz] );
    e[1] = 0.25*(double)( m[emy[v]].epsy*ey[v]*ey[v] + m[emy[vz]].epsy*ey[vz]*ey[vz] + m[emy[vx]].epsy*ey[vx]*ey[vx] + m[emy[vzx]].epsy*ey[vzx]*ey[vzx] );
    e[2] = 0.25*(double)( m[emz[v]].epsz*ez[v]*ez[v] + m[emz[vx]].epsz*ez[vx]*ez[vx] + m[emz[vy]].epsz*ez[vy]*ez[vy] + m[emz[vxy]].epsz*ez[vxy]*ez[vxy] );
    e[3] = 0.5 *(double)( m[fmx[v]].rmux*bx[v]*bx[v] + m[fmx[vx]].rmux*bx[vx]*bx[vx] );
    e[4] = 0.5 *(double)( m[fmy[v]].rmuy*by[v]*by[v] + m[fmy[vy]].rmuy*by[vy]*by[vy] );
    e[5] = 0.5 *(double)( m[fmz[v]].rmuz*bz[v]*bz[v] + m[fmz[vz]].rmuz*bz[vz]*bz[vz] );
  }
# pragma unroll
  for( int c=0; c<6; c++ ) val[ (size_t)c*ncell + i ] = e[c];
}

__global__ void __launch_bounds__(256)
k_energy_p_values( double * __restrict__ val, const vpb_particle_t * __restrict__ p, int np,
                   const float * __restrict__ ip, float qdt_2mc, float msp ) {
  const int n = blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  const float one = 1.0f;
  const float4 * pp = reinterpret_cast<const float4 *>( p + n );
  const float4 r0 = __ldg( pp ), r1 = __ldg( pp+1 );
  const float dx = r0.x, dy = r0.y, dz = r0.z;
  const Interp f = load_interp( ip, __float_as_int( r0.w ) );
  float v0 = r1.x + qdt_2mc*( ( f.ex + dy*f.dexdy ) + dz*( f.dexdz + dy*f.d2exdydz ) );
  float v1 = r1.y + qdt_2mc*( ( f.ey + dz*f.deydz ) + dx*( f.deydx + dz*f.d2eydzdx ) );
  float v2 = r1.z + qdt_2mc*( ( f.ez + dx*f.dezdx ) + dy*( f.dezdy + dx*f.d2ezdxdy ) );
  v0 = ( v0*v0 + v1*v1 ) + v2*v2;
  v0 = ( msp*r1.w )*( v0/( one + sqrtf( one + v0 ) ) );
  val[n] = (double)v0;
}

// thread (r,c): the share of pipeline r (r==P: the host's remainder) of component c
__global__ void k_ordered_shares( double * __restrict__ share, const double * __restrict__ val, int n, int ncomp, int P ) {
  const int t = blockIdx.x*blockDim.x + threadIdx.x;
  if( t>=( P+1 )*ncomp ) return;
  const int r = t/ncomp, c = t - r*ncomp;
  const double tt = (double)( n/16 )/(double)P;
  const int i0 = 16*(int)( tt*(double)r + 0.5 );
  const int i1 = r<P ? 16*(int)( tt*(double)( r+1 ) + 0.5 ) : n;
  const double * v = val + (size_t)c*n;
  double e = 0.0;
  for( int i=i0; i<i1; i++ ) e += v[i];
  share[t] = e;
}

__global__ void k_ordered_total( double * __restrict__ out, const double * __restrict__ share, int ncomp, int P, double scale ) {
  const int c = threadIdx.x;
  if( c>=ncomp ) return;
  double tot = 0.0;
  for( int r=0; r<=P; r++ ) tot += share[ r*ncomp + c ];
  out[c] = tot*scale;
}

namespace {
struct ParBuf { void * p = nullptr; size_t bytes = 0; };
ParBuf B_cnt, B_off, B_bin, B_val, B_perm, B_start, B_big, B_dbl;
void * need( ParBuf & b, size_t bytes ) {
  if( bytes>b.bytes ) {
    VPB_CUDA( cudaStreamSynchronize( rt().stream ) );
    if( b.p ) VPB_CUDA( cudaFree( b.p ) );
    b.bytes = bytes + bytes/8 + 4096;
    VPB_CUDA( cudaMalloc( &b.p, b.bytes ) );
  }
  return b.p;
}
}

void release_parity_buffers() {
  for( ParBuf * b : { &B_cnt, &B_off, &B_bin, &B_val, &B_perm, &B_start, &B_big, &B_dbl } ) { if( b->p ) cudaFree( b->p ); b->p = nullptr; b->bytes = 0; }
}

static void ensure_private_copies( AccMirror * ma, int P ) {
  Runtime & r = rt();
  if( ma->n_priv==P && ma->d_priv ) return;
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( ma->d_priv ) VPB_CUDA( cudaFree( ma->d_priv ) );
  const size_t bytes = sizeof(float)*ACC_STRIDE*(size_t)ma->grid->dev.nv*(size_t)P;
  VPB_CUDA( cudaMalloc( &ma->d_priv, bytes ) );
  VPB_CUDA( cudaMemsetAsync( ma->d_priv, 0, bytes, r.stream ) );
  ma->n_priv = P;
}

void advance_p_parity( SpeciesMirror * ms, AccMirror * ma, InterpMirror * mi, const PushCoef & c ) {
  Runtime & r = rt();
  const int P = ma->host->n_pipeline;
  if( P<1 || P>9 ) VPB_ERROR( "PARITY deposit reproduces the reference's order for 1..9 pipelines (aa->n_pipeline=%d)", P );
  if( ms->np>( 1<<28 ) ) VPB_ERROR( "PARITY deposit is a verification mode for up to 2^28 particles per species" );
  ensure_private_copies( ma, P );
  const GridDev & g = ms->grid->dev;
  const int np = ms->np;
  ParRanges R; R.P = P;
  const double t = (double)( np/16 )/(double)P;                  // DISTRIBUTE, src/util/util_base.h:191-197
  for( int k=0; k<=P; k++ ) R.start[k] = 16*(int)( t*(double)k + 0.5 );
  R.start[P+1] = np;
  int * cnt = (int *)need( B_cnt, sizeof(int)*( (size_t)np+32 ) );
  int * off = (int *)need( B_off, sizeof(int)*( (size_t)np+32 ) );
  int * tiles = (int *)scratch( sizeof(int)*( (size_t)div_up( np, 4096 ) + 64 ) );
  k_parity_push<true><<< div_up(np,128), 128, 0, r.stream >>>( ms->d_p, np, mi->d, cnt, nullptr, nullptr, nullptr, nullptr, 0, nullptr, g, c, R );
  exclusive_scan( cnt, off, nullptr, tiles, np );
  VPB_CHECK_LAUNCH(); r.launches++;
  int last[2];
  d2h( last+0, off+np-1, sizeof(int), r.stream );
  d2h( last+1, cnt+np-1, sizeof(int), r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  const int nrec = last[0] + last[1];
  int   * rec_bin = (int *)need( B_bin, sizeof(int)*( (size_t)nrec+32 ) );
  float * rec_val = (float *)need( B_val, sizeof(float)*12*( (size_t)nrec+8 ) );
  k_parity_push<false><<< div_up(np,128), 128, 0, r.stream >>>( ms->d_p, np, mi->d, cnt, off, rec_bin, rec_val,
                                                               ms->d_pm, (int)ms->cap_pm, ms->d_counters, g, c, R );
  VPB_CHECK_LAUNCH(); r.launches++;
  const int nbin = ( P+1 )*g.nv;
  int * start = (int *)need( B_start, sizeof(int)*( (size_t)nbin+64 ) );
  int * perm  = (int *)need( B_perm, sizeof(int)*( (size_t)nrec+32 ) );
  int * big   = (int *)need( B_big, sizeof(int)*( 2*(size_t)nrec+64 ) );
  bin_indices( rec_bin, nrec, nbin, start, perm );
  k_parity_sum<<< div_up( nbin, FIX_WARPS ), 32*FIX_WARPS, 0, r.stream >>>( start, nbin, nrec, perm, big, rec_val, ma->d, ma->d_priv, g.nv );
  VPB_CHECK_LAUNCH(); r.launches++;
}

// ---- accumulate_rho_p in the reference's order (rho_p.cc:22-112) -------------
// The reference adds each particle's eight node charges in particle order, so
// per node the sum runs over contributing (particle, corner) records in
// ascending record index.  Records of a chunk of particles are binned by node,
// each node's list is put in order by one warp and added sequentially; chunks
// follow one another in particle order, so the result is bit-identical.
__global__ void __launch_bounds__(256)
k_rho_records( int * __restrict__ key, float * __restrict__ val, const vpb_particle_t * __restrict__ p, int n,
               int sy, int sz, float q_8V ) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if( i>=n ) return;
  float w0 = p[i].dx, w1 = p[i].dy, w2, w3, w4, w5, w6, w7 = p[i].w*q_8V; const float dz = p[i].dz;
  const int v = p[i].i;
  w6 = w7 - w0*w7; w7 = w7 + w0*w7;
  w4 = w6 - w1*w6; w5 = w7 - w1*w7;
  w6 = w6 + w1*w6; w7 = w7 + w1*w7;
  w0 = w4 - dz*w4; w1 = w5 - dz*w5; w2 = w6 - dz*w6; w3 = w7 - dz*w7;
  w4 = w4 + dz*w4; w5 = w5 + dz*w5; w6 = w6 + dz*w6; w7 = w7 + dz*w7;
  int * k = key + 8*(size_t)i; float * o = val + 8*(size_t)i;
  k[0] = v;    k[1] = v+1;    k[2] = v+sy;    k[3] = v+sy+1;
  k[4] = v+sz; k[5] = v+sz+1; k[6] = v+sz+sy; k[7] = v+sz+sy+1;
  o[0] = w0; o[1] = w1; o[2] = w2; o[3] = w3; o[4] = w4; o[5] = w5; o[6] = w6; o[7] = w7;
}

__global__ void __launch_bounds__(32*FIX_WARPS)
k_rho_sum( const int * __restrict__ start, int nbin, int nrec, int * __restrict__ perm, int * __restrict__ big_scratch,
           const float * __restrict__ rec_val, float * __restrict__ rhof ) {
  __shared__ int s[FIX_WARPS][FIX_SMEM];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int bin = blockIdx.x*FIX_WARPS + w;
  if( bin>=nbin ) return;
  const int b = start[bin], e = bin+1<nbin ? start[bin+1] : nrec, n = e-b;
  if( n<=0 ) return;
  int n2 = 32; while( n2<n ) n2 <<= 1;
  int * a = n2<=FIX_SMEM ? s[w] : big_scratch + 2*(size_t)b;
  for( int t=lane; t<n2; t+=32 ) a[t] = t<n ? perm[b+t] : 0x7fffffff;
  __threadfence_block(); __syncwarp();
  bitonic_mem( a, n2, lane );
  if( lane==0 ) {
    float v = rhof[bin];
    for( int t=0; t<n; t++ ) v = v + rec_val[ a[t] ];
    rhof[bin] = v;
  }
}

void accumulate_rho_p_parity( FieldMirror * mf, SpeciesMirror * ms, float q_8V ) {
  Runtime & r = rt();
  const GridDev & g = mf->grid->dev;
  const int CHUNK = 1<<22;                            // particles per pass (32 M records)
  for( long long first=0; first<ms->np; first+=CHUNK ) {
    const int n = ms->np-first<CHUNK ? (int)( ms->np-first ) : CHUNK, nrec = 8*n;
    int   * key = (int *)need( B_bin, sizeof(int)*( (size_t)nrec+32 ) );
    float * val = (float *)need( B_val, sizeof(float)*( (size_t)nrec+32 ) );
    k_rho_records<<< div_up( n, 256 ), 256, 0, r.stream >>>( key, val, ms->d_p + first, n, g.sy, g.sz, q_8V );
    VPB_CHECK_LAUNCH(); r.launches++;
    int * start = (int *)need( B_start, sizeof(int)*( (size_t)g.nv+64 ) );
    int * perm  = (int *)need( B_perm, sizeof(int)*( (size_t)nrec+32 ) );
    int * big   = (int *)need( B_big, sizeof(int)*( 2*(size_t)nrec+64 ) );
    bin_indices( key, nrec, g.nv, start, perm );
    k_rho_sum<<< div_up( g.nv, FIX_WARPS ), 32*FIX_WARPS, 0, r.stream >>>( start, g.nv, nrec, perm, big, val, mf->comp( F_RHOF ) );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
}

// ---- accumulate_hydro_p in the reference's order (hydro_p.c:62-165) ----------
// Same scheme as accumulate_rho_p above with 14 moments per record: the
// per-particle part (half Boris step, velocities, weights) is computed once
// into a 16-float record; every node's (particle, corner) records are put in
// ascending order by one warp, whose lanes form the 14 increments of 32
// records at a time and add them one record after the other.
__global__ void __launch_bounds__(256)
k_hydro_records( int * __restrict__ key, float * __restrict__ rec, const vpb_particle_t * __restrict__ p, int n,
                 const float * __restrict__ ip, int sy, int sz, HydroCoef k ) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if( i>=n ) return;
  const float4 * pp = reinterpret_cast<const float4 *>( p + i );
  const float4 r0 = __ldg( pp ), r1 = __ldg( pp+1 );
  const int v = __float_as_int( r0.w );
  const HydroParticle h = hydro_particle( r0, r1, load_interp( ip, v ), k );
  int * kk = key + 8*(size_t)i;
  kk[0] = v;    kk[1] = v+1;    kk[2] = v+sy;    kk[3] = v+sy+1;
  kk[4] = v+sz; kk[5] = v+sz+1; kk[6] = v+sz+sy; kk[7] = v+sz+sy+1;
  float4 * o = reinterpret_cast<float4 *>( rec + 16*(size_t)i );
  o[0] = make_float4( h.wn[0], h.wn[1], h.wn[2], h.wn[3] );
  o[1] = make_float4( h.wn[4], h.wn[5], h.wn[6], h.wn[7] );
  o[2] = make_float4( h.ux, h.uy, h.uz, h.vx );
  o[3] = make_float4( h.vy, h.vz, h.ke_mc, 0.f );
}

__global__ void __launch_bounds__(32*FIX_WARPS)
k_hydro_sum( const int * __restrict__ start, int nbin, int nrec, int * __restrict__ perm, int * __restrict__ big_scratch,
             const float * __restrict__ rec, float * __restrict__ hy, HydroCoef k ) {
  __shared__ int s[FIX_WARPS][FIX_SMEM];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int bin = blockIdx.x*FIX_WARPS + w;
  if( bin>=nbin ) return;
  const int b = start[bin], e = bin+1<nbin ? start[bin+1] : nrec, n = e-b;
  if( n<=0 ) return;
  int n2 = 32; while( n2<n ) n2 <<= 1;
  int * a = n2<=FIX_SMEM ? s[w] : big_scratch + 2*(size_t)b;
  for( int t=lane; t<n2; t+=32 ) a[t] = t<n ? perm[b+t] : 0x7fffffff;
  __threadfence_block(); __syncwarp();
  bitonic_mem( a, n2, lane );
  float * h = hy + (size_t)bin*HY_STRIDE;
  float acc[14];
# pragma unroll
  for( int c=0; c<14; c++ ) acc[c] = h[c];
  for( int t0=0; t0<n; t0+=32 ) {
    const int cnt = n-t0<32 ? n-t0 : 32;
    float o[14];
#   pragma unroll
    for( int c=0; c<14; c++ ) o[c] = 0.f;
    if( lane<cnt ) {
      const int r = a[t0+lane];
      const float * q = rec + 16*(size_t)( r>>3 );
      hydro_node_terms( o, q[ r&7 ], k, q[8], q[9], q[10], q[11], q[12], q[13], q[14] );
    }
    for( int t=0; t<cnt; t++ ) {
#     pragma unroll
      for( int c=0; c<14; c++ ) acc[c] = acc[c] + __shfl_sync( 0xffffffffu, o[c], t );
    }
  }
  if( lane==0 ) {
#   pragma unroll
    for( int c=0; c<14; c++ ) h[c] = acc[c];
  }
}

void accumulate_hydro_p_parity( HydroMirror * mh, SpeciesMirror * ms, InterpMirror * mi, const HydroCoef & k ) {
  Runtime & r = rt();
  const GridDev & g = mh->grid->dev;
  const int CHUNK = 1<<22;                            // particles per pass (32 M records)
  for( long long first=0; first<ms->np; first+=CHUNK ) {
    const int n = ms->np-first<CHUNK ? (int)( ms->np-first ) : CHUNK, nrec = 8*n;
    int   * key = (int *)need( B_bin, sizeof(int)*( (size_t)nrec+32 ) );
    float * rec = (float *)need( B_val, sizeof(float)*16*( (size_t)n+8 ) );
    k_hydro_records<<< div_up( n, 256 ), 256, 0, r.stream >>>( key, rec, ms->d_p + first, n, mi->d, g.sy, g.sz, k );
    VPB_CHECK_LAUNCH(); r.launches++;
    int * start = (int *)need( B_start, sizeof(int)*( (size_t)g.nv+64 ) );
    int * perm  = (int *)need( B_perm, sizeof(int)*( (size_t)nrec+32 ) );
    int * big   = (int *)need( B_big, sizeof(int)*( 2*(size_t)nrec+64 ) );
    bin_indices( key, nrec, g.nv, start, perm );
    k_hydro_sum<<< div_up( g.nv, FIX_WARPS ), 32*FIX_WARPS, 0, r.stream >>>( start, g.nv, nrec, perm, big, rec, mh->d, k );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
}

void reduce_accumulators_parity( AccMirror * ma ) {
  Runtime & r = rt();
  if( !ma->d_priv || ma->n_priv<1 ) return;
  const size_t n = (size_t)ACC_STRIDE*(size_t)ma->grid->dev.nv;
  k_parity_reduce<<< div_up( (long long)n, 256 ), 256, 0, r.stream >>>( ma->d, ma->d_priv, ma->n_priv, n );
  VPB_CHECK_LAUNCH(); r.launches++;
}

static void ordered_sum( double * host_out, const double * val, int n, int ncomp, double scale ) {
  Runtime & r = rt();
  const int P = r.ref_pipelines;
  double * share = (double *)scratch( sizeof(double)*( (size_t)( P+1 )*ncomp + ncomp + 8 ) );
  k_ordered_shares<<< div_up( ( P+1 )*ncomp, 32 ), 32, 0, r.stream >>>( share, val, n, ncomp, P );
  k_ordered_total<<< 1, 32, 0, r.stream >>>( share + ( P+1 )*ncomp, share, ncomp, P, scale );
  VPB_CHECK_LAUNCH(); r.launches += 2;
  d2h( host_out, share + ( P+1 )*ncomp, sizeof(double)*ncomp, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
}

void energy_f_parity( double * en6, FieldMirror * m ) {
  Runtime & r = rt();
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = m->host->g;
  const int ncell = g.nx*g.ny*g.nz;
  double * val = (double *)need( B_dbl, sizeof(double)*6*(size_t)ncell );
  if( m->n_mc==1 ) k_energy_f_values<true ><<< div_up(ncell,256), 256, 0, r.stream >>>( val, m->d, m->d_mat, m->d_mc, g );
  else             k_energy_f_values<false><<< div_up(ncell,256), 256, 0, r.stream >>>( val, m->d, m->d_mat, m->d_mc, g );
  VPB_CHECK_LAUNCH(); r.launches++;
  ordered_sum( en6, val, ncell, 6, 0.5*hg->eps0*hg->dV );
}

double energy_p_parity( SpeciesMirror * ms, InterpMirror * mi, float qdt_2mc, float msp ) {
  Runtime & r = rt();
  const int np = ms->np;
  double * val = (double *)need( B_dbl, sizeof(double)*( (size_t)np+8 ) );
  if( np>0 ) { k_energy_p_values<<< div_up(np,256), 256, 0, r.stream >>>( val, ms->d_p, np, mi->d, qdt_2mc, msp ); r.launches++; }
  VPB_CHECK_LAUNCH();
  double local = 0;
  ordered_sum( &local, val, np, 1, 1.0 );
  return local;
}

} // namespace vpb

using namespace vpb;

extern "C" void vpic_b200_set_reference_pipelines( int n_pipeline ) {
  if( n_pipeline<1 || n_pipeline>PAR_MAXP ) VPB_ERROR( "Bad args" );
  rt().ref_pipelines = n_pipeline;
}
