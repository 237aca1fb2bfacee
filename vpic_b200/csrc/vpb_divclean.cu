// vpb_divclean.cu -- the interval-gated field kernels of the reference's
// field_advance_kernels_t (src/field_advance/field_advance.h:170-218):
//   clear_rhof / accumulate_rho_p / synchronize_rho          (rho_p.cc:22-112, remote.c:534-624)
//   compute_div_e_err / compute_rms_div_e_err / clean_div_e   ({vacuum_,}*_pipeline.{h,c})
//   compute_div_b_err / compute_rms_div_b_err / clean_div_b   (*_pipeline.{h,c,cc})
//   synchronize_tang_e_norm_b                                 (remote.c:298-415)
//   compute_rhob / compute_curl_b                             (initialisation, {vacuum_,}*_pipeline)
// and the local / remote ghost and adjust steps they use (local.c:128-222,
// 298-445; remote.c:137-296).
//
// All are HBM-bound streaming stencils over the SoA field planes: one thread
// per lattice point on the flat plane-ordered grid of k_advance_b (blocks walk
// the z planes in order, neighbour planes hit L2).  Every per-point expression
// keeps the reference's operation order (this file is compiled -fmad=false),
// so every field they write is bit-identical to the reference's scalar build.
// The two double-precision diagnostics are summed in a fixed order
// (deterministic) but not in the reference's order: 1e-12 relative.
#include "vpb_runtime.h"

namespace vpb {

namespace {

inline dim3 lattice_grid( const GridDev & g, int ex ) {      // (nx+ex)(ny+ex) x (nz+ex) points
  return dim3( div_up( (long long)(g.nx+ex)*(g.ny+ex), 256 ), g.nz+ex );
}

// ---- div E error / bound charge -------------------------------------------
// compute_div_e_err: 36 B read (ex,ey,ez as 3 planes + neighbours from L2,
// rhof, rhob) + 4 B written per node; compute_rhob likewise.
struct DivECoef { float px, py, pz, cj, nc; };

template<bool RHOB, bool VACUUM>
__global__ void __launch_bounds__(256)
k_div_e( float * __restrict__ f, const int16_t * __restrict__ mat, const vpb_material_coefficient_t * __restrict__ m,
         GridDev g, DivECoef c ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  const float * __restrict__ ex = f + F_EX*P, * __restrict__ ey = f + F_EY*P, * __restrict__ ez = f + F_EZ*P;
  const float * __restrict__ rhof = f + F_RHOF*P;
  float * __restrict__ rhob = f + F_RHOB*P, * __restrict__ derr = f + F_DIVE*P;
  const int wx = w-1, wy = w-g.sy, wz = w-g.sz;
  float d, nc;
  if( VACUUM ) {
    d = c.px*( ex[w] - ex[wx] ) + c.py*( ey[w] - ey[wy] ) + c.pz*( ez[w] - ez[wz] );
    nc = c.nc;
  } else {
    const int16_t * emx = mat + 0*P, * emy = mat + 1*P, * emz = mat + 2*P, * nm = mat + 3*P;
    d = c.px*( m[emx[w]].epsx*ex[w] - m[emx[wx]].epsx*ex[wx] ) +
        c.py*( m[emy[w]].epsy*ey[w] - m[emy[wy]].epsy*ey[wy] ) +
        c.pz*( m[emz[w]].epsz*ez[w] - m[emz[wz]].epsz*ez[wz] );
    nc = m[nm[w]].nonconductive;
  }
  if( RHOB ) rhob[w] = nc*( d - rhof[w] );
  else       derr[w] = nc*( d - c.cj*( rhof[w] + rhob[w] ) );
}

// ---- Marder pass on E -------------------------------------------------------
template<bool VACUUM>
__global__ void __launch_bounds__(256)
k_clean_div_e( float * __restrict__ f, const int16_t * __restrict__ mat, const vpb_material_coefficient_t * __restrict__ m,
               GridDev g, float px, float py, float pz ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  float * __restrict__ ex = f + F_EX*P, * __restrict__ ey = f + F_EY*P, * __restrict__ ez = f + F_EZ*P;
  const float * __restrict__ derr = f + F_DIVE*P;
  const float d0 = derr[w];
  if( VACUUM ) {
    if( x<g.nx ) ex[w] += px*( derr[w+1]    - d0 );
    if( y<g.ny ) ey[w] += py*( derr[w+g.sy] - d0 );
    if( z<g.nz ) ez[w] += pz*( derr[w+g.sz] - d0 );
  } else {
    const int16_t * emx = mat + 0*P, * emy = mat + 1*P, * emz = mat + 2*P;
    if( x<g.nx ) ex[w] += m[emx[w]].drivex*px*( derr[w+1]    - d0 );
    if( y<g.ny ) ey[w] += m[emy[w]].drivey*py*( derr[w+g.sy] - d0 );
    if( z<g.nz ) ez[w] += m[emz[w]].drivez*pz*( derr[w+g.sz] - d0 );
  }
}

// ---- div B error and Marder pass on B ---------------------------------------
__global__ void __launch_bounds__(256)
k_div_b( float * __restrict__ f, GridDev g, float px, float py, float pz ) {
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=g.nx*g.ny ) return;
  const int z = blockIdx.y, y = ixy / g.nx, x = ixy - y*g.nx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  const float * __restrict__ cbx = f + F_CBX*P, * __restrict__ cby = f + F_CBY*P, * __restrict__ cbz = f + F_CBZ*P;
  ( f + F_DIVB*P )[w] = px*( cbx[w+1] - cbx[w] ) + py*( cby[w+g.sy] - cby[w] ) + pz*( cbz[w+g.sz] - cbz[w] );
}

__global__ void __launch_bounds__(256)
k_clean_div_b( float * __restrict__ f, GridDev g, float px, float py, float pz ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  float * __restrict__ cbx = f + F_CBX*P, * __restrict__ cby = f + F_CBY*P, * __restrict__ cbz = f + F_CBZ*P;
  const float * __restrict__ derr = f + F_DIVB*P;
  const bool inx = x<g.nx, iny = y<g.ny, inz = z<g.nz;
  const float d0 = derr[w];
  if( iny && inz ) cbx[w] += px*( d0 - derr[w-1]    );
  if( inz && inx ) cby[w] += py*( d0 - derr[w-g.sy] );
  if( inx && iny ) cbz[w] += pz*( d0 - derr[w-g.sz] );
}

// ---- tca = curl( cB / mu_r ) -------------------------------------------------
struct CurlCoef { float px_muz, px_muy, py_mux, py_muz, pz_muy, pz_mux; };

template<bool VACUUM>
__global__ void __launch_bounds__(256)
k_curl_b( float * __restrict__ f, const int16_t * __restrict__ mat, const vpb_material_coefficient_t * __restrict__ m,
          GridDev g, CurlCoef c, float px, float py, float pz ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const int vx = v-1, vy = v-g.sy, vz = v-g.sz;
  const size_t P = (size_t)g.nvp;
  const float * __restrict__ cbx = f + F_CBX*P, * __restrict__ cby = f + F_CBY*P, * __restrict__ cbz = f + F_CBZ*P;
  float * __restrict__ tx = f + F_TCAX*P, * __restrict__ ty = f + F_TCAY*P, * __restrict__ tz = f + F_TCAZ*P;
  if( VACUUM ) {
    if( x<g.nx ) tx[v] = ( c.py_muz*( cbz[v] - cbz[vy] ) - c.pz_muy*( cby[v] - cby[vz] ) );
    if( y<g.ny ) ty[v] = ( c.pz_mux*( cbx[v] - cbx[vz] ) - c.px_muz*( cbz[v] - cbz[vx] ) );
    if( z<g.nz ) tz[v] = ( c.px_muy*( cby[v] - cby[vx] ) - c.py_mux*( cbx[v] - cbx[vy] ) );
  } else {
    const int16_t * fmx = mat + 4*P, * fmy = mat + 5*P, * fmz = mat + 6*P;
    if( x<g.nx ) tx[v] = ( py*( cbz[v]*m[fmz[v]].rmuz - cbz[vy]*m[fmz[vy]].rmuz ) -
                           pz*( cby[v]*m[fmy[v]].rmuy - cby[vz]*m[fmy[vz]].rmuy ) );
    if( y<g.ny ) ty[v] = ( pz*( cbx[v]*m[fmx[v]].rmux - cbx[vz]*m[fmx[vz]].rmux ) -
                           px*( cbz[v]*m[fmz[v]].rmuz - cbz[vx]*m[fmz[vx]].rmuz ) );
    if( z<g.nz ) tz[v] = ( px*( cby[v]*m[fmy[v]].rmuy - cby[vx]*m[fmy[vx]].rmuy ) -
                           py*( cbx[v]*m[fmx[v]].rmux - cbx[vy]*m[fmx[vy]].rmux ) );
  }
}

// ---- rms diagnostics ---------------------------------------------------------
// Per-block double partials over a fixed grid, then one block adds them in
// index order: run-to-run deterministic.
constexpr int RMS_BLOCK = 256;

__device__ __forceinline__ void block_sum_to( double v, double * __restrict__ partial ) {
  __shared__ double s[RMS_BLOCK/32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for( int o=16; o; o>>=1 ) v += __shfl_down_sync( 0xffffffffu, v, o );
  if( lane==0 ) s[w] = v;
  __syncthreads();
  if( threadIdx.x==0 ) {
    double t = 0;
    for( int k=0; k<RMS_BLOCK/32; k++ ) t += s[k];
    partial[blockIdx.x] = t;
  }
}

// compute_rms_div_e_err_pipeline.c:53-166: nodes (1:nx+1,1:ny+1,1:nz+1);
// interior nodes enter as FLOAT squares, boundary nodes as double products
// weighted 1/2 per boundary axis.
__global__ void __launch_bounds__(RMS_BLOCK)
k_rms_div_e( double * __restrict__ partial, const float * __restrict__ f, GridDev g ) {
  const int mx = g.nx+1, my = g.ny+1, mz = g.nz+1;
  const long long n = (long long)mx*my*mz;
  const float * __restrict__ derr = f + F_DIVE*(size_t)g.nvp;
  double e = 0;
  for( long long i = blockIdx.x*(long long)blockDim.x + threadIdx.x; i<n; i += (long long)gridDim.x*blockDim.x ) {
    const int x = (int)( i % mx ), y = (int)( ( i / mx ) % my ), z = (int)( i / ( (long long)mx*my ) );
    const float d = derr[ (x+1) + g.sy*(y+1) + g.sz*(z+1) ];
    const int nb = ( x==0 || x==g.nx ) + ( y==0 || y==g.ny ) + ( z==0 || z==g.nz );
    if( nb==0 ) e += (double)( d*d );
    else e += ( nb==1 ? 0.5 : nb==2 ? 0.25 : 0.125 )*(double)d*(double)d;
  }
  block_sum_to( e, partial );
}

// compute_rms_div_b_err_pipeline.c:7-103: cells (1:nx,1:ny,1:nz), float squares
__global__ void __launch_bounds__(RMS_BLOCK)
k_rms_div_b( double * __restrict__ partial, const float * __restrict__ f, GridDev g ) {
  const long long n = (long long)g.nx*g.ny*g.nz;
  const float * __restrict__ derr = f + F_DIVB*(size_t)g.nvp;
  double e = 0;
  for( long long i = blockIdx.x*(long long)blockDim.x + threadIdx.x; i<n; i += (long long)gridDim.x*blockDim.x ) {
    const int x = (int)( i % g.nx ), y = (int)( ( i / g.nx ) % g.ny ), z = (int)( i / ( (long long)g.nx*g.ny ) );
    const float d = derr[ (x+1) + g.sy*(y+1) + g.sz*(z+1) ];
    e += (double)( d*d );
  }
  block_sum_to( e, partial );
}

__global__ void k_sum_ordered( double * __restrict__ out, const double * __restrict__ partial, int n ) {
  // one block: thread t adds partial[t], partial[t+T], ... ; fixed tree on top
  __shared__ double s[RMS_BLOCK];
  double v = 0;
  for( int i=threadIdx.x; i<n; i+=RMS_BLOCK ) v += partial[i];
  s[threadIdx.x] = v;
  __syncthreads();
  for( int o=RMS_BLOCK/2; o; o>>=1 ) { if( threadIdx.x<o ) s[threadIdx.x] += s[threadIdx.x+o]; __syncthreads(); }
  if( threadIdx.x==0 ) out[0] = s[0];
}

// ---- accumulate_rho_p (rho_p.cc:22-112) --------------------------------------
// 32 B read per particle; each thread walks RP_K consecutive particles and keeps
// the 8 node sums of the current voxel in registers (cell-sorted particles
// stay in one voxel for most of a run), committing with 8 float atomics on a
// voxel change.  The reference adds in particle order, so rhof agrees to
// rounding (FAST-deposit tolerance), not bit for bit.
constexpr int RP_K = 8;

__device__ __forceinline__ void commit_rho( float * __restrict__ rhof, int v, int sy, int sz, float * __restrict__ s ) {
  atomicAdd( rhof + v,         s[0] ); atomicAdd( rhof + v + 1,         s[1] );
  atomicAdd( rhof + v + sy,    s[2] ); atomicAdd( rhof + v + sy + 1,    s[3] );
  atomicAdd( rhof + v + sz,    s[4] ); atomicAdd( rhof + v + sz + 1,    s[5] );
  atomicAdd( rhof + v + sz+sy, s[6] ); atomicAdd( rhof + v + sz+sy + 1, s[7] );
# pragma unroll
  for( int k=0; k<8; k++ ) s[k] = 0.f;
}

__global__ void __launch_bounds__(256)
k_accumulate_rho_p( float * __restrict__ rhof, const vpb_particle_t * __restrict__ p, int np, int sy, int sz, float q_8V ) {
  const long long first = ( blockIdx.x*(long long)blockDim.x + threadIdx.x )*RP_K;
  if( first>=np ) return;
  const int cnt = np-first<RP_K ? (int)( np-first ) : RP_K;
  float s[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
  int cur = -1;
  for( int k=0; k<cnt; k++ ) {
    const float4 r0 = reinterpret_cast<const float4 *>( p + first + k )[0];
    const float wq  = p[first+k].w;
    const int v = __float_as_int( r0.w );
    if( v!=cur ) { if( cur>=0 ) commit_rho( rhof, cur, sy, sz, s ); cur = v; }
    float w0 = r0.x, w1 = r0.y, w2, w3, w4, w5, w6, w7 = wq*q_8V; const float dz = r0.z;
    w6 = w7 - w0*w7; w7 = w7 + w0*w7;
    w4 = w6 - w1*w6; w5 = w7 - w1*w7;
    w6 = w6 + w1*w6; w7 = w7 + w1*w7;
    w0 = w4 - dz*w4; w1 = w5 - dz*w5; w2 = w6 - dz*w6; w3 = w7 - dz*w7;
    w4 = w4 + dz*w4; w5 = w5 + dz*w5; w6 = w6 + dz*w6; w7 = w7 + dz*w7;
    s[0] += w0; s[1] += w1; s[2] += w2; s[3] += w3; s[4] += w4; s[5] += w5; s[6] += w6; s[7] += w7;
  }
  if( cur>=0 ) commit_rho( rhof, cur, sy, sz, s );
}

// ---- ghost / boundary planes of this family ----------------------------------
// Index sets on the plane normal to X (remote.c:16-41), u along Y, w along Z:
//   NODE: u 1..nY+1, w 1..nZ+1     FACE: u 1..nY, w 1..nZ
//   YZ edge (component Y): u 1..nY, w 1..nZ+1    ZY edge (component Z): u 1..nY+1, w 1..nZ
enum NodeOpCode {
  NO_NE_GHOST,      // local.c:128-179  ghost eX,tcaX from the interior (mode: +1 copy, -1 negate, 2 extrapolate)
  NO_NE_PACK,       // remote.c:154-164 eX of a node plane -> buf
  NO_NE_UNPACK,     // remote.c:181-190 ghost eX = rw*buf + lw*interior
  NO_DB_GHOST,      // local.c:181-222  ghost div_b_err (mode: +1, -1, 0)
  NO_DB_PACK, NO_DB_UNPACK,            // remote.c:226-236, 253-263
  NO_DERR_ZERO,     // local.c:298-333  div_e_err = 0 on the face nodes
  NO_RHO_ADJUST,    // local.c:376-445  mode 0: rhof = 0, 1: rhof *= 2, 2: rhob = 0
  NO_RHO_PACK, NO_RHO_FOLD,            // remote.c:552-587
  NO_SYNC_PACK, NO_SYNC_AVG            // remote.c:316-375
};

struct NodeArgs {
  int op, X, side, plane, mode;
  float w0, w1, w2, w3;      // weights
  float * buf;               // packed message
  double * sq;               // NO_SYNC_AVG: squared mismatch per item
};

__global__ void __launch_bounds__(128)
k_node_op( float * __restrict__ f, GridDev g, NodeArgs a ) {
  const int n3[3] = { g.nx, g.ny, g.nz };
  const int s3[3] = { 1, g.sy, g.sz };
  const int X = a.X, Y = (X+1)%3, Z = (X+2)%3;
  const int nY = n3[Y], nZ = n3[Z], sX = s3[X], sY = s3[Y], sZ = s3[Z];
  const size_t P = (size_t)g.nvp;
  const int t = blockIdx.x*blockDim.x + threadIdx.x;
  const bool face_set = ( a.op==NO_DB_GHOST || a.op==NO_DB_PACK || a.op==NO_DB_UNPACK );
  if( a.op==NO_SYNC_PACK || a.op==NO_SYNC_AVG ) {
    // three item sets one after the other: FACE (cbX), YZ edges (eY,tcaY), ZY edges (eZ,tcaZ)
    const int nF = nY*nZ, nA = nY*(nZ+1), nB = (nY+1)*nZ;
    if( t>=nF+nA+nB ) return;
    int k, ue, comp, slot; bool pair;
    if( t<nF )         { k = t;       ue = nY;   comp = F_CBX+X; slot = t;                   pair = false; }
    else if( t<nF+nA ) { k = t-nF;    ue = nY;   comp = F_EX+Y;  slot = nF + 2*k;            pair = true;  }
    else               { k = t-nF-nA; ue = nY+1; comp = F_EX+Z;  slot = nF + 2*nA + 2*k;     pair = true;  }
    const int v = a.plane*sX + ( 1 + k%ue )*sY + ( 1 + k/ue )*sZ;
    float * c0 = f + (size_t)comp*P;
    float * c1 = pair ? f + (size_t)( comp - F_EX + F_TCAX )*P : nullptr;
    if( a.op==NO_SYNC_PACK ) {
      a.buf[slot] = c0[v];
      if( pair ) a.buf[slot+1] = c1[v];
    } else {
      double w1 = a.buf[slot], w2 = c0[v];
      c0[v] = 0.5*( w1+w2 );
      a.sq[t] = ( w1-w2 )*( w1-w2 );
      if( pair ) { w1 = a.buf[slot+1]; w2 = c1[v]; c1[v] = 0.5*( w1+w2 ); }
    }
    return;
  }
  const int ue = face_set ? nY : nY+1, we = face_set ? nZ : nZ+1;
  if( t>=ue*we ) return;
  const int v = a.plane*sX + ( 1 + t%ue )*sY + ( 1 + t/ue )*sZ;
  switch( a.op ) {
  case NO_NE_GHOST: {
    float * e = f + (size_t)( F_EX+X )*P, * tc = f + (size_t)( F_TCAX+X )*P;
    const int v1 = v - a.side*sX, v2 = v - 2*a.side*sX;
    if( a.mode==2 ) { e[v] = 2*e[v1] - e[v2]; tc[v] = 2*tc[v1] - tc[v2]; }
    else if( a.mode<0 ) { e[v] = -e[v1]; tc[v] = -tc[v1]; }
    else { e[v] = e[v1]; tc[v] = tc[v1]; }
  } break;
  case NO_NE_PACK: a.buf[t] = f[ (size_t)( F_EX+X )*P + v ]; break;
  case NO_NE_UNPACK: { float * e = f + (size_t)( F_EX+X )*P; e[v] = a.w0*a.buf[t] + a.w1*e[ v - a.side*sX ]; } break;
  case NO_DB_GHOST: {
    float * d = f + (size_t)F_DIVB*P;
    const float s = d[ v - a.side*sX ];
    d[v] = a.mode==0 ? 0.f : ( a.mode<0 ? -s : s );
  } break;
  case NO_DB_PACK: a.buf[t] = f[ (size_t)F_DIVB*P + v ]; break;
  case NO_DB_UNPACK: { float * d = f + (size_t)F_DIVB*P; d[v] = a.w0*a.buf[t] + a.w1*d[ v - a.side*sX ]; } break;
  case NO_DERR_ZERO: f[ (size_t)F_DIVE*P + v ] = 0.f; break;
  case NO_RHO_ADJUST:
    if( a.mode==0 ) f[ (size_t)F_RHOF*P + v ] = 0.f;
    else if( a.mode==1 ) f[ (size_t)F_RHOF*P + v ] *= 2;
    else f[ (size_t)F_RHOB*P + v ] = 0.f;
    break;
  case NO_RHO_PACK: a.buf[2*t] = f[ (size_t)F_RHOF*P + v ]; a.buf[2*t+1] = f[ (size_t)F_RHOB*P + v ]; break;
  case NO_RHO_FOLD: {
    float * rf = f + (size_t)F_RHOF*P, * rb = f + (size_t)F_RHOB*P;
    rf[v] = a.w0*rf[v] + a.w1*a.buf[2*t];          // lw, rw
    rb[v] = a.w2*rb[v] + a.w3*a.buf[2*t+1];        // hlw, hrw
  } break;
  }
}

int node_items( const GridDev & g, int X, int op ) {
  const int n3[3] = { g.nx, g.ny, g.nz };
  const int nY = n3[(X+1)%3], nZ = n3[(X+2)%3];
  if( op==NO_SYNC_PACK || op==NO_SYNC_AVG ) return nY*nZ + nY*(nZ+1) + (nY+1)*nZ;
  if( op==NO_DB_GHOST || op==NO_DB_PACK || op==NO_DB_UNPACK ) return nY*nZ;
  return (nY+1)*(nZ+1);
}

void launch_node( FieldMirror * m, NodeArgs a ) {
  Runtime & r = rt();
  const GridDev & g = m->grid->dev;
  const int n = node_items( g, a.X, a.op );
  if( n<=0 ) return;
  k_node_op<<< div_up( n, 128 ), 128, 0, r.stream >>>( m->d, g, a );
  VPB_CHECK_LAUNCH(); r.launches++;
}

inline bool is_local( int bc ) { return bc<0 || bc>=g_world_size; }

// Ghost planes of one scalar/normal quantity: local boundary conditions, then
// the remote (or periodic-to-self) neighbours, axis by axis.
// kind 0: normal E (nodes; message = eX of planes 1 / nX), kind 1: div_b_err (faces).
void fill_ghost_planes( FieldMirror * m, int kind ) {
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev; const vpb_grid_t * hg = gm->host;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const float d3[3] = { hg->dx, hg->dy, hg->dz };
  const int op_ghost = kind==0 ? NO_NE_GHOST : NO_DB_GHOST, op_pack = kind==0 ? NO_NE_PACK : NO_DB_PACK,
            op_unpack = kind==0 ? NO_NE_UNPACK : NO_DB_UNPACK;
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( gm, X, side );
    if( !is_local( bc ) ) continue;
    NodeArgs a{}; a.op = op_ghost; a.X = X; a.side = side; a.plane = side<0 ? 0 : n3[X]+1;
    switch( bc ) {
    case VPB_PEC_FIELDS:       a.mode = 1; break;
    case VPB_SYMMETRIC_FIELDS:
    case VPB_PMC_FIELDS:       a.mode = -1; break;
    case VPB_ABSORB_FIELDS:    a.mode = kind==0 ? 2 : 0; break;
    default: VPB_ERROR( "Bad boundary condition encountered." );
    }
    launch_node( m, a );
  }
  for( int X=0; X<3; X++ ) {
    const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
    if( !rem_lo && !rem_hi ) continue;
    const int nf = node_items( g, X, op_pack );
    if( rem_lo ) { NodeArgs a{}; a.op = op_pack; a.X = X; a.side = -1; a.plane = 1;     a.buf = m->d_face;                  launch_node( m, a ); }
    if( rem_hi ) { NodeArgs a{}; a.op = op_pack; a.X = X; a.side = +1; a.plane = n3[X]; a.buf = m->d_face + m->face_floats; launch_node( m, a ); }
    float * from_lo, * from_hi;
    exchange_axis( m, X, nf, &from_lo, &from_hi );
    float lw = d3[X];                                  // equal cells on both sides
    const float rw = (2.*d3[X])/(lw+d3[X]);
    lw = (lw-d3[X])/(lw+d3[X]);
    if( from_lo ) { NodeArgs a{}; a.op = op_unpack; a.X = X; a.side = -1; a.plane = 0;       a.w0 = rw; a.w1 = lw; a.buf = from_lo; launch_node( m, a ); }
    if( from_hi ) { NodeArgs a{}; a.op = op_unpack; a.X = X; a.side = +1; a.plane = n3[X]+1; a.w0 = rw; a.w1 = lw; a.buf = from_hi; launch_node( m, a ); }
  }
}

// local_adjust_rhof / local_adjust_rhob (local.c:376-445); which: 0 rhof, 1 rhob
void local_adjust_rho( FieldMirror * m, int which ) {
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( gm, X, side );
    if( !is_local( bc ) ) continue;
    NodeArgs a{}; a.op = NO_RHO_ADJUST; a.X = X; a.side = side; a.plane = side<0 ? 1 : n3[X]+1;
    switch( bc ) {
    case VPB_PEC_FIELDS: a.mode = which ? 2 : 0; launch_node( m, a ); break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS: case VPB_ABSORB_FIELDS:
      if( !which ) { a.mode = 1; launch_node( m, a ); }
      break;
    default: VPB_ERROR( "Bad boundary condition encountered." );
    }
  }
}

// scratch layout for the double diagnostics: [0] result, [8..8+n) partials
double sum_partials_to_host( double * base, int n ) {
  Runtime & r = rt();
  k_sum_ordered<<< 1, RMS_BLOCK, 0, r.stream >>>( base, base+8, n );
  VPB_CHECK_LAUNCH(); r.launches++;
  double h = 0;
  d2h( &h, base, sizeof(double), r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  return h;
}

int rms_blocks( long long n ) {
  long long nb = ( n + RMS_BLOCK - 1 )/RMS_BLOCK;
  const long long cap = (long long)rt().sm_count*8;
  return (int)( nb<1 ? 1 : nb>cap ? cap : nb );
}

FieldMirror * field_on_device( const vpb_field_array_t * fa ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  FieldMirror * m = mirror( const_cast<vpb_field_array_t *>( fa ) );
  to_device( m );
  return m;
}

} // namespace

} // namespace vpb

using namespace vpb;

extern "C" {

// rho_p.cc:22-112
void accumulate_rho_p( vpb_field_array_t * fa, const vpb_species_t * sp ) {
  VPB_TRACE_CALL( "accumulate_rho_p" );
  if( !fa || !sp || fa->g!=sp->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  SpeciesMirror * ms = mirror( const_cast<vpb_species_t *>( sp ) );
  to_device( ms );
  const GridDev & g = m->grid->dev;
  if( ms->np>0 && r.deposit_mode==VPB_DEPOSIT_PARITY ) accumulate_rho_p_parity( m, ms, sp->q*sp->g->r8V );
  else if( ms->np>0 ) {
    const long long nthread = ( (long long)ms->np + RP_K - 1 )/RP_K;
    k_accumulate_rho_p<<< div_up( nthread, 256 ), 256, 0, r.stream >>>( m->comp( F_RHOF ), ms->d_p, ms->np, g.sy, g.sz, sp->q*sp->g->r8V );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
  finish_write( m );
}

// remote.c:534-624
void vpic_b200_synchronize_rho( vpb_field_array_t * fa ) {
  FieldMirror * m = field_on_device( fa );
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev; const vpb_grid_t * hg = gm->host;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const float d3[3] = { hg->dx, hg->dy, hg->dz };
  local_adjust_rho( m, 0 );
  local_adjust_rho( m, 1 );
  for( int X=0; X<3; X++ ) {
    const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
    if( !rem_lo && !rem_hi ) continue;
    const int nf = 2*node_items( g, X, NO_RHO_PACK );
    if( rem_lo ) { NodeArgs a{}; a.op = NO_RHO_PACK; a.X = X; a.side = -1; a.plane = 1;       a.buf = m->d_face;                  launch_node( m, a ); }
    if( rem_hi ) { NodeArgs a{}; a.op = NO_RHO_PACK; a.X = X; a.side = +1; a.plane = n3[X]+1; a.buf = m->d_face + m->face_floats; launch_node( m, a ); }
    float * from_lo, * from_hi;
    exchange_axis( m, X, nf, &from_lo, &from_hi );
    float hrw = d3[X], hlw = hrw + d3[X];              // equal cells on both sides
    hrw /= hlw; hlw = d3[X]/hlw;
    const float lw = hlw + hlw, rw = hrw + hrw;
    if( from_lo ) { NodeArgs a{}; a.op = NO_RHO_FOLD; a.X = X; a.side = -1; a.plane = 1;       a.w0 = lw; a.w1 = rw; a.w2 = hlw; a.w3 = hrw; a.buf = from_lo; launch_node( m, a ); }
    if( from_hi ) { NodeArgs a{}; a.op = NO_RHO_FOLD; a.X = X; a.side = +1; a.plane = n3[X]+1; a.w0 = lw; a.w1 = rw; a.w2 = hlw; a.w3 = hrw; a.buf = from_hi; launch_node( m, a ); }
  }
  finish_write( m );
}

static void div_e_stencil( vpb_field_array_t * fa, FieldMirror * m, bool want_rhob ) {
  Runtime & r = rt();
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * mc = prm->mc;
  fill_ghost_planes( m, 0 );
  const bool vacuum = m->n_mc==1;
  DivECoef c; c.cj = 1./hg->eps0; c.nc = mc->nonconductive;
  if( want_rhob ) {
    if( vacuum ) { c.px = (g.nx>1) ? hg->eps0*mc->epsx*hg->rdx : 0; c.py = (g.ny>1) ? hg->eps0*mc->epsy*hg->rdy : 0; c.pz = (g.nz>1) ? hg->eps0*mc->epsz*hg->rdz : 0; }
    else         { c.px = (g.nx>1) ? hg->eps0*hg->rdx : 0;          c.py = (g.ny>1) ? hg->eps0*hg->rdy : 0;          c.pz = (g.nz>1) ? hg->eps0*hg->rdz : 0; }
  } else {
    if( vacuum ) { c.px = ((g.nx>1) ? hg->rdx : 0)*mc->epsx; c.py = ((g.ny>1) ? hg->rdy : 0)*mc->epsy; c.pz = ((g.nz>1) ? hg->rdz : 0)*mc->epsz; }
    else         { c.px = (g.nx>1) ? hg->rdx : 0;            c.py = (g.ny>1) ? hg->rdy : 0;            c.pz = (g.nz>1) ? hg->rdz : 0; }
  }
  const dim3 grid = lattice_grid( g, 1 );
  if( want_rhob ) {
    if( vacuum ) k_div_e<true, true ><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c );
    else         k_div_e<true, false><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c );
  } else {
    if( vacuum ) k_div_e<false,true ><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c );
    else         k_div_e<false,false><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c );
  }
  VPB_CHECK_LAUNCH(); r.launches++;
}

// {vacuum_,}compute_div_e_err_pipeline.c + local_adjust_div_e (local.c:298-333)
void vpic_b200_compute_div_e_err( vpb_field_array_t * fa ) {
  FieldMirror * m = field_on_device( fa );
  div_e_stencil( fa, m, false );
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( gm, X, side );
    if( !is_local( bc ) ) continue;
    if( bc==VPB_PEC_FIELDS || bc==VPB_ABSORB_FIELDS ) {
      NodeArgs a{}; a.op = NO_DERR_ZERO; a.X = X; a.side = side; a.plane = side<0 ? 1 : n3[X]+1;
      launch_node( m, a );
    } else if( bc!=VPB_SYMMETRIC_FIELDS && bc!=VPB_PMC_FIELDS ) VPB_ERROR( "Bad boundary condition encountered." );
  }
  finish_write( m );
}

// {vacuum_,}compute_rhob_pipeline.c + local_adjust_rhob
void vpic_b200_compute_rhob( vpb_field_array_t * fa ) {
  FieldMirror * m = field_on_device( fa );
  div_e_stencil( fa, m, true );
  local_adjust_rho( m, 1 );
  finish_write( m );
}

// compute_rms_div_e_err_pipeline.c:53-166
double vpic_b200_compute_rms_div_e_err( const vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const int nb = rms_blocks( (long long)(g.nx+1)*(g.ny+1)*(g.nz+1) );
  double * base = (double *)scratch( sizeof(double)*( 8 + (size_t)nb ) );
  k_rms_div_e<<< nb, RMS_BLOCK, 0, r.stream >>>( base+8, m->d, g );
  VPB_CHECK_LAUNCH(); r.launches++;
  double local[2] = { sum_partials_to_host( base, nb )*hg->dV, ( hg->nx*hg->ny*hg->nz )*hg->dV };
  if( g_world_size>1 ) comm_allsum_d( local, 2 );
  return hg->eps0*sqrt( local[0]/local[1] );
}

// {vacuum_,}clean_div_e_pipeline.c + local_adjust_tang_e
void vpic_b200_clean_div_e( vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const vpb_material_coefficient_t * mc = ( (const vpb_sfa_params_t *)fa->params )->mc;
  const float _rdx = (g.nx>1) ? hg->rdx : 0, _rdy = (g.ny>1) ? hg->rdy : 0, _rdz = (g.nz>1) ? hg->rdz : 0;
  const float alphadt = 0.3888889/( _rdx*_rdx + _rdy*_rdy + _rdz*_rdz );
  const dim3 grid = lattice_grid( g, 1 );
  if( m->n_mc==1 ) {
    const float px = (alphadt*_rdx)*mc->drivex, py = (alphadt*_rdy)*mc->drivey, pz = (alphadt*_rdz)*mc->drivez;
    k_clean_div_e<true ><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, px, py, pz );
  } else {
    const float px = alphadt*_rdx, py = alphadt*_rdy, pz = alphadt*_rdz;
    k_clean_div_e<false><<< grid, 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, px, py, pz );
  }
  VPB_CHECK_LAUNCH(); r.launches++;
  local_adjust_tang_e( m );
  finish_write( m );
}

// compute_div_b_err_pipeline.cc
void vpic_b200_compute_div_b_err( vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const float px = (g.nx>1) ? hg->rdx : 0, py = (g.ny>1) ? hg->rdy : 0, pz = (g.nz>1) ? hg->rdz : 0;
  k_div_b<<< dim3( div_up( (long long)g.nx*g.ny, 256 ), g.nz ), 256, 0, r.stream >>>( m->d, g, px, py, pz );
  VPB_CHECK_LAUNCH(); r.launches++;
  finish_write( m );
}

// compute_rms_div_b_err_pipeline.c
double vpic_b200_compute_rms_div_b_err( const vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const int nb = rms_blocks( (long long)g.nx*g.ny*g.nz );
  double * base = (double *)scratch( sizeof(double)*( 8 + (size_t)nb ) );
  k_rms_div_b<<< nb, RMS_BLOCK, 0, r.stream >>>( base+8, m->d, g );
  VPB_CHECK_LAUNCH(); r.launches++;
  double local[2] = { sum_partials_to_host( base, nb )*hg->dV, ( hg->nx*hg->ny*hg->nz )*hg->dV };
  if( g_world_size>1 ) comm_allsum_d( local, 2 );
  return hg->eps0*sqrt( local[0]/local[1] );
}

// clean_div_b_pipeline.cc + local_adjust_norm_b
void vpic_b200_clean_div_b( vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  float px = (g.nx>1) ? hg->rdx : 0, py = (g.ny>1) ? hg->rdy : 0, pz = (g.nz>1) ? hg->rdz : 0;
  const float alphadt = 0.3888889/( px*px + py*py + pz*pz );
  px *= alphadt; py *= alphadt; pz *= alphadt;
  fill_ghost_planes( m, 1 );
  k_clean_div_b<<< lattice_grid( g, 1 ), 256, 0, r.stream >>>( m->d, g, px, py, pz );
  VPB_CHECK_LAUNCH(); r.launches++;
  local_adjust_norm_b( m );
  finish_write( m );
}

// {vacuum_,}compute_curl_b_pipeline.cc
void vpic_b200_compute_curl_b( vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const vpb_material_coefficient_t * mc = ( (const vpb_sfa_params_t *)fa->params )->mc;
  fill_ghost_tang_b( m );
  const float px = (g.nx>1) ? hg->cvac*hg->dt*hg->rdx : 0, py = (g.ny>1) ? hg->cvac*hg->dt*hg->rdy : 0,
              pz = (g.nz>1) ? hg->cvac*hg->dt*hg->rdz : 0;
  CurlCoef c;
  c.px_muz = px*mc->rmuz; c.px_muy = px*mc->rmuy; c.py_mux = py*mc->rmux;
  c.py_muz = py*mc->rmuz; c.pz_muy = pz*mc->rmuy; c.pz_mux = pz*mc->rmux;
  if( m->n_mc==1 ) k_curl_b<true ><<< lattice_grid( g, 1 ), 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c, px, py, pz );
  else             k_curl_b<false><<< lattice_grid( g, 1 ), 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, c, px, py, pz );
  VPB_CHECK_LAUNCH(); r.launches++;
  local_adjust_tang_e( m );
  finish_write( m );
}

// remote.c:298-415
double vpic_b200_synchronize_tang_e_norm_b( vpb_field_array_t * fa ) {
  Runtime & r = rt();
  FieldMirror * m = field_on_device( fa );
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  local_adjust_tang_e( m );
  local_adjust_norm_b( m );
  double err = 0;
  for( int X=0; X<3; X++ ) {
    const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
    if( !rem_lo && !rem_hi ) continue;
    const int ni = node_items( g, X, NO_SYNC_PACK );
    const int n3Y = n3[(X+1)%3], n3Z = n3[(X+2)%3];
    const int nf = 2*n3Y*(n3Z+1) + 2*n3Z*(n3Y+1) + n3Y*n3Z;
    if( rem_lo ) { NodeArgs a{}; a.op = NO_SYNC_PACK; a.X = X; a.side = -1; a.plane = 1;       a.buf = m->d_face;                  launch_node( m, a ); }
    if( rem_hi ) { NodeArgs a{}; a.op = NO_SYNC_PACK; a.X = X; a.side = +1; a.plane = n3[X]+1; a.buf = m->d_face + m->face_floats; launch_node( m, a ); }
    float * from_lo, * from_hi;
    exchange_axis( m, X, nf, &from_lo, &from_hi );
    double * base = (double *)scratch( sizeof(double)*( 8 + 2*(size_t)ni ) );
    int nsq = 0;
    // face nX+1 first, like the reference's END_RECV order
    if( from_hi ) { NodeArgs a{}; a.op = NO_SYNC_AVG; a.X = X; a.side = +1; a.plane = n3[X]+1; a.buf = from_hi; a.sq = base+8+nsq; launch_node( m, a ); nsq += ni; }
    if( from_lo ) { NodeArgs a{}; a.op = NO_SYNC_AVG; a.X = X; a.side = -1; a.plane = 1;       a.buf = from_lo; a.sq = base+8+nsq; launch_node( m, a ); nsq += ni; }
    if( nsq ) err += sum_partials_to_host( base, nsq );
  }
  finish_write( m );
  if( g_world_size>1 ) comm_allsum_d( &err, 1 );
  return err;
}

} // extern "C"
