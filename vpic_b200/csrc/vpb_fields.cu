// vpb_fields.cu -- Yee FDTD kernels on the device SoA field planes.
//   advance_b   reference: src/field_advance/standard/pipeline/advance_b_pipeline.{h,cc}
//   advance_e   reference: .../pipeline/{vacuum_,}advance_e_pipeline.{h,cc}
//   energy_f    reference: .../pipeline/{vacuum_,}energy_f_pipeline.{h,c}
//   ghost / boundary planes: src/field_advance/standard/{local,remote}.c
// Compiled with -fmad=false: every expression rounds like the reference.
#include "vpb_runtime.h"
#include <cstdio>

namespace vpb {

// ---------------------------------------------------------------------------
// Bulk stencils.  One thread per lattice point of the (nx+1)(ny+1)(nz+1) block
// that holds every updated component; x fastest => unit-stride plane access.
// advance_b: 24 B read + 12 B written per cell = 36 B.
// ---------------------------------------------------------------------------

// Lattice point i (x fastest over the (nx+1)(ny+1)(nz+1) block) -> voxel.
__device__ __forceinline__ int lattice_voxel( const GridDev & g, int i, int mx, int my, int & x, int & y, int & z ) {
  x = i % mx; const int t = i / mx; y = t % my; z = t / my;
  return (x+1) + g.sy*(y+1) + g.sz*(z+1);
}

// One thread per lattice point, blockIdx.y = z: one integer division per
// thread, no grid-stride loop (tools/membench.cu: a flat grid streams 6R+3W
// float planes at 6.5 TB/s, a persistent one at 6.0).  Out-of-range neighbours
// are clamped instead of branched so all twelve loads issue together.
__global__ void __launch_bounds__(256)
k_advance_b( float * __restrict__ f, GridDev g, float px, float py, float pz ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  const float * __restrict__ ex = f + F_EX*P, * __restrict__ ey = f + F_EY*P, * __restrict__ ez = f + F_EZ*P;
  float * __restrict__ cbx = f + F_CBX*P, * __restrict__ cby = f + F_CBY*P, * __restrict__ cbz = f + F_CBZ*P;
  const bool inx = x<g.nx, iny = y<g.ny, inz = z<g.nz;
  const int wx = inx ? w+1 : w, wy = iny ? w+g.sy : w, wz = inz ? w+g.sz : w;
  const float ex0 = ex[w], ey0 = ey[w], ez0 = ez[w];
  const float ezy = ez[wy], eyz = ey[wz], exz = ex[wz], ezx = ez[wx], eyx = ey[wx], exy = ex[wy];
  const float bx = cbx[w], by = cby[w], bz = cbz[w];
  if( iny && inz ) cbx[w] = bx - ( py*( ezy-ez0 ) - pz*( eyz-ey0 ) );
  if( inz && inx ) cby[w] = by - ( pz*( exz-ex0 ) - px*( ezx-ez0 ) );
  if( inx && iny ) cbz[w] = bz - ( px*( eyx-ey0 ) - py*( exy-ex0 ) );
}

// advance_e, vacuum variant (one material): 48 B read + 24 B written per cell.
struct VacuumCoef {
  float decayx, drivex, decayy, drivey, decayz, drivez;
  float px_muz, px_muy, py_mux, py_muz, pz_muy, pz_mux, damp, cj;
};

__global__ void __launch_bounds__(256)
k_advance_e_vacuum( float * __restrict__ f, GridDev g, VacuumCoef c ) {
  const int mx = g.nx+1, my = g.ny+1;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  if( ixy>=mx*my ) return;
  const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
  const int w = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const size_t P = (size_t)g.nvp;
  const float * __restrict__ cbx = f + F_CBX*P, * __restrict__ cby = f + F_CBY*P, * __restrict__ cbz = f + F_CBZ*P;
  const float * __restrict__ jfx = f + F_JFX*P, * __restrict__ jfy = f + F_JFY*P, * __restrict__ jfz = f + F_JFZ*P;
  float * __restrict__ ex = f + F_EX*P, * __restrict__ ey = f + F_EY*P, * __restrict__ ez = f + F_EZ*P;
  float * __restrict__ tx = f + F_TCAX*P, * __restrict__ ty = f + F_TCAY*P, * __restrict__ tz = f + F_TCAZ*P;
  // the -x,-y,-z neighbours always exist (ghost layer): all loads issue together
  const float bx0 = cbx[w], by0 = cby[w], bz0 = cbz[w];
  const float bzy = cbz[w-g.sy], byz = cby[w-g.sz], bxz = cbx[w-g.sz], bzx = cbz[w-1], byx = cby[w-1], bxy = cbx[w-g.sy];
  const float e0 = ex[w], e1 = ey[w], e2 = ez[w], t0 = tx[w], t1 = ty[w], t2 = tz[w], j0 = jfx[w], j1 = jfy[w], j2 = jfz[w];
  if( x<g.nx ) {
    const float tca = ( c.py_muz*( bz0 - bzy ) - c.pz_muy*( by0 - byz ) ) - c.damp*t0;
    tx[w] = tca;
    ex[w] = c.decayx*e0 + c.drivex*( tca - c.cj*j0 );
  }
  if( y<g.ny ) {
    const float tca = ( c.pz_mux*( bx0 - bxz ) - c.px_muz*( bz0 - bzx ) ) - c.damp*t1;
    ty[w] = tca;
    ey[w] = c.decayy*e1 + c.drivey*( tca - c.cj*j1 );
  }
  if( z<g.nz ) {
    const float tca = ( c.px_muy*( by0 - byx ) - c.py_mux*( bx0 - bxy ) ) - c.damp*t2;
    tz[w] = tca;
    ez[w] = c.decayz*e2 + c.drivez*( tca - c.cj*j2 );
  }
}

// advance_e with per-point materials (advance_e_pipeline.h:50-75).
// mat planes: 0 ematx 1 ematy 2 ematz 3 nmat 4 fmatx 5 fmaty 6 fmatz 7 cmat.
__global__ void __launch_bounds__(256)
k_advance_e_material( float * __restrict__ f, const int16_t * __restrict__ mat,
                      const vpb_material_coefficient_t * __restrict__ m, GridDev g,
                      float px, float py, float pz, float damp, float cj ) {
  const int mx = g.nx+1, my = g.ny+1, mz = g.nz+1;
  const int n = mx*my*mz;
  const size_t P = (size_t)g.nvp;
  const float * cbx = f + F_CBX*P, * cby = f + F_CBY*P, * cbz = f + F_CBZ*P;
  const float * jfx = f + F_JFX*P, * jfy = f + F_JFY*P, * jfz = f + F_JFZ*P;
  float * ex = f + F_EX*P, * ey = f + F_EY*P, * ez = f + F_EZ*P;
  float * tx = f + F_TCAX*P, * ty = f + F_TCAY*P, * tz = f + F_TCAZ*P;
  const int16_t * emx = mat + 0*P, * emy = mat + 1*P, * emz = mat + 2*P;
  const int16_t * fmx = mat + 4*P, * fmy = mat + 5*P, * fmz = mat + 6*P;
  {
    const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
    if( ixy>=mx*my ) return;
    const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
    (void)n; (void)mz;
    const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
    const int vx = v-1, vy = v-g.sy, vz = v-g.sz;
    if( x<g.nx ) {
      const float tca = ( py*( cbz[v]*m[fmz[v]].rmuz - cbz[vy]*m[fmz[vy]].rmuz ) -
                          pz*( cby[v]*m[fmy[v]].rmuy - cby[vz]*m[fmy[vz]].rmuy ) ) - damp*tx[v];
      tx[v] = tca;
      ex[v] = m[emx[v]].decayx*ex[v] + m[emx[v]].drivex*( tca - cj*jfx[v] );
    }
    if( y<g.ny ) {
      const float tca = ( pz*( cbx[v]*m[fmx[v]].rmux - cbx[vz]*m[fmx[vz]].rmux ) -
                          px*( cbz[v]*m[fmz[v]].rmuz - cbz[vx]*m[fmz[vx]].rmuz ) ) - damp*ty[v];
      ty[v] = tca;
      ey[v] = m[emy[v]].decayy*ey[v] + m[emy[v]].drivey*( tca - cj*jfy[v] );
    }
    if( z<g.nz ) {
      const float tca = ( px*( cby[v]*m[fmy[v]].rmuy - cby[vx]*m[fmy[vx]].rmuy ) -
                          py*( cbx[v]*m[fmx[v]].rmux - cbx[vy]*m[fmx[vy]].rmux ) ) - damp*tz[v];
      tz[v] = tca;
      ez[v] = m[emz[v]].decayz*ez[v] + m[emz[v]].drivez*( tca - cj*jfz[v] );
    }
  }
}

// ---------------------------------------------------------------------------
// energy_f: float products, double sums.  Per-block partials in a fixed grid,
// then one block adds them in index order => run-to-run deterministic.
// ---------------------------------------------------------------------------

constexpr int EN_BLOCK = 256;

template<bool VACUUM>
__global__ void __launch_bounds__(EN_BLOCK)
k_energy_f( double * __restrict__ partial, const float * __restrict__ f, const int16_t * __restrict__ mat,
            const vpb_material_coefficient_t * __restrict__ m, GridDev g ) {
  const int ncell = g.nx*g.ny*g.nz;
  const size_t P = (size_t)g.nvp;
  const float * ex = f + F_EX*P, * ey = f + F_EY*P, * ez = f + F_EZ*P;
  const float * bx = f + F_CBX*P, * by = f + F_CBY*P, * bz = f + F_CBZ*P;
  double e[6] = {0,0,0,0,0,0};
  float qepsx = 0, qepsy = 0, qepsz = 0, hrmux = 0, hrmuy = 0, hrmuz = 0;
  if( VACUUM ) {
    qepsx = 0.25*m->epsx; qepsy = 0.25*m->epsy; qepsz = 0.25*m->epsz;
    hrmux = 0.50*m->rmux; hrmuy = 0.50*m->rmuy; hrmuz = 0.50*m->rmuz;
  }
  for( int i = blockIdx.x*blockDim.x + threadIdx.x; i<ncell; i += gridDim.x*blockDim.x ) {
    const int x = i % g.nx, t = i / g.nx, y = t % g.ny, z = t / g.ny;
    const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
    const int vx = v+1, vy = v+g.sy, vz = v+g.sz, vyz = v+g.sy+g.sz, vzx = v+g.sz+1, vxy = v+1+g.sy;
    if( VACUUM ) {
      e[0] += (double)( qepsx*( ex[v]*ex[v] + ex[vy]*ex[vy] + ex[vz]*ex[vz] + ex[vyz]*ex[vyz] ) );
      e[1] += (double)( qepsy*( ey[v]*ey[v] + ey[vz]*ey[vz] + ey[vx]*ey[vx] + ey[vzx]*ey[vzx] ) );
      e[2] += (double)( qepsz*( ez[v]*ez[v] + ez[vx]*ez[vx] + ez[vy]*ez[vy] + ez[vxy]*ez[vxy] ) );
      e[3] += (double)( hrmux*( bx[v]*bx[v] + bx[vx]*bx[vx] ) );
      e[4] += (double)( hrmuy*( by[v]*by[v] + by[vy]*by[vy] ) );
      e[5] += (double)( hrmuz*( bz[v]*bz[v] + bz[vz]*bz[vz] ) );
    } else {
      const int16_t * emx = mat + 0*P, * emy = mat + 1*P, * emz = mat + 2*P;
      const int16_t * fmx = mat + 4*P, * fmy = mat + 5*P, * fmz = mat + 6*P;
      e[0] += 0.25*(double)( m[emx[v]].epsx*ex[v]*ex[v] + m[emx[vy]].epsx*ex[vy]*ex[vy] + m[emx[vz]].epsx*ex[vz]*ex[vz] + m[emx[vyz]].epsx*ex[vyz]*ex[vyz] );
      e[1] += 0.25*(double)( m[emy[v]].epsy*ey[v]*ey[v] + m[emy[vz]].epsy*ey[vz]*ey[vz] + m[emy[vx]].epsy*ey[vx]*ey[vx] + m[emy[vzx]].epsy*ey[vzx]*ey[vzx] );
      e[2] += 0.25*(double)( m[emz[v]].epsz*ez[v]*ez[v] + m[emz[vx]].epsz*ez[vx]*ez[vx] + m[emz[vy]].epsz*ez[vy]*ez[vy] + m[emz[vxy]].epsz*ez[vxy]*ez[vxy] );
      e[3] += 0.5 *(double)( m[fmx[v]].rmux*bx[v]*bx[v] + m[fmx[vx]].rmux*bx[vx]*bx[vx] );
      e[4] += 0.5 *(double)( m[fmy[v]].rmuy*by[v]*by[v] + m[fmy[vy]].rmuy*by[vy]*by[vy] );
      e[5] += 0.5 *(double)( m[fmz[v]].rmuz*bz[v]*bz[v] + m[fmz[vz]].rmuz*bz[vz]*bz[vz] );
    }
  }
  __shared__ double s[6][EN_BLOCK/32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
# pragma unroll
  for( int c=0; c<6; c++ ) {
    double vsum = e[c];
    for( int o=16; o; o>>=1 ) vsum += __shfl_down_sync( 0xffffffffu, vsum, o );
    if( lane==0 ) s[c][w] = vsum;
  }
  __syncthreads();
  if( threadIdx.x<6 ) {
    double vsum = 0;
    for( int k=0; k<EN_BLOCK/32; k++ ) vsum += s[threadIdx.x][k];
    partial[ (size_t)blockIdx.x*6 + threadIdx.x ] = vsum;
  }
}

__global__ void k_sum_partials( double * __restrict__ out, const double * __restrict__ partial, int nblock, int ncomp, double scale ) {
  const int c = threadIdx.x;
  if( c>=ncomp ) return;
  double s = 0;
  for( int b=0; b<nblock; b++ ) s += partial[ (size_t)b*ncomp + c ];
  out[c] = s*scale;
}

// ---------------------------------------------------------------------------
// Boundary / ghost planes.  Every operation of local.c / remote.c works on the
// plane normal to axis X (Y=(X+1)%3, Z=(X+2)%3) over two index sets:
//   "B-kind" (tangential B):  comp Y on u:1..nY+1,w:1..nZ ; comp Z on u:1..nY,w:1..nZ+1
//   "E-kind" (tangential E/J): comp Y on u:1..nY,w:1..nZ+1 ; comp Z on u:1..nY+1,w:1..nZ
// (remote.c:16-41).  One thread per plane item; these are tiny launches.
// ---------------------------------------------------------------------------

enum PlaneOpCode {
  OP_PACK_B, OP_UNPACK_GHOST_B, OP_PACK_J, OP_FOLD_J,
  OP_GHOST_B_COPY,      // PEC: ghost = +interior ; symmetric/pmc: ghost = -interior  (sign in w0)
  OP_GHOST_B_ABSORB,    // Higdon ABC (local.c:80-107)
  OP_TANG_E_ZERO,       // PEC: e = tca = 0 on the face (local.c:237-247)
  OP_JF_SCALE,          // PEC: jf = 0 ; others: jf *= 2 (local.c:345-354); factor in w0
  OP_NORM_B_ZERO        // symmetric: cbX = 0 on the face (local.c:282-284)
};

struct PlaneArgs {
  int op, X, side;       // side: -1 low face, +1 high face
  int plane;             // plane index the items live on (ghost / face)
  float w0, w1;          // weights (rw,lw) / sign / factor / (decay, drive)
  float cdt_dX, cdt_dY, cdt_dZ;
  float * buf;           // packed face buffer (a local one, or a mailbox slot: the neighbour's for a pack, ours for an unpack)
  // fused exchange (vpb_comm.cu): sync = 1: after the last block has stored its part of buf in the
  // neighbour's mailbox, publish seq in *flag; sync = 2: wait until *flag has reached seq before reading buf
  int sync; unsigned seq; unsigned * flag; unsigned * done;
};

__device__ __forceinline__ void plane_wait( const PlaneArgs & a ) {
  if( a.sync==2 ) {
    if( threadIdx.x==0 ) {
      unsigned v; const long long t0 = clock64();
      do {
        asm volatile( "ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flag) : "memory" );
        if( (int)( v-a.seq )>=0 ) break;
        if( clock64()-t0>1200000000000ll ) { printf( "[vpic_b200] fused exchange: plane %u never arrived (flag %u)\n", a.seq, v ); __trap(); }
        __nanosleep( 64 );
      } while( true );
    }
    __syncthreads();
  }
}
__device__ __forceinline__ void plane_publish( const PlaneArgs & a, unsigned * counter ) {
  if( a.sync==1 ) {
    __threadfence_system();
    __syncthreads();
    if( threadIdx.x==0 && atomicAdd( counter, 1u )==gridDim.x-1 ) {
      *counter = 0;
      asm volatile( "st.release.sys.global.u32 [%0], %1;" :: "l"(a.flag), "r"(a.seq) : "memory" );
    }
  }
}

__device__ __forceinline__ void plane_op_item( float * __restrict__ f, const GridDev & g, const PlaneArgs & a, const int t ) {
  const int n3[3] = { g.nx, g.ny, g.nz };
  const int s3[3] = { 1, g.sy, g.sz };
  const int X = a.X, Y = (X+1)%3, Z = (X+2)%3;
  const int nY = n3[Y], nZ = n3[Z], sX = s3[X], sY = s3[Y], sZ = s3[Z];
  const size_t P = (size_t)g.nvp;
  const bool bkind = ( a.op==OP_PACK_B || a.op==OP_UNPACK_GHOST_B || a.op==OP_GHOST_B_COPY || a.op==OP_GHOST_B_ABSORB );
  int nA, uA, nB, uB;                       // item counts and u-extents of the two sets
  if( a.op==OP_NORM_B_ZERO ) { uA = nY; nA = nY*nZ; uB = 1; nB = 0; }
  else if( bkind ) { uA = nY+1; nA = (nY+1)*nZ; uB = nY;   nB = nY*(nZ+1); }
  else             { uA = nY;   nA = nY*(nZ+1); uB = nY+1; nB = (nY+1)*nZ; }
  if( t>=nA+nB ) return;
  const bool isA = t<nA;
  const int k = isA ? t : t-nA;
  const int ue = isA ? uA : uB;
  const int u = 1 + k%ue, w = 1 + k/ue;
  const int v = a.plane*sX + u*sY + w*sZ;
  const int comp = isA ? Y : Z;             // tangential component index
  switch( a.op ) {
  case OP_PACK_B: a.buf[t] = f[ (F_CBX+comp)*P + v ]; break;
  case OP_PACK_J: a.buf[t] = f[ (F_JFX+comp)*P + v ]; break;
  case OP_UNPACK_GHOST_B: {                 // remote.c:108-115: ghost = rw*recv + lw*interior
    float * cb = f + (F_CBX+comp)*P;
    cb[v] = a.w0*a.buf[t] + a.w1*cb[ v - a.side*sX ];
  } break;
  case OP_FOLD_J: {                         // remote.c:463-470: jf = lw*jf + rw*recv
    float * jf = f + (F_JFX+comp)*P;
    jf[v] = a.w1*jf[v] + a.w0*a.buf[t];
  } break;
  case OP_GHOST_B_COPY: {
    float * cb = f + (F_CBX+comp)*P;
    const float s = cb[ v - a.side*sX ];
    cb[v] = a.w0<0 ? -s : s;
  } break;
  case OP_GHOST_B_ABSORB: {
    // a.plane is the ghost plane; face plane = ghost for the high side, 1 for the low side
    const int vf = ( a.side<0 ? 1 : a.plane )*sX + u*sY + w*sZ;
    const int fh = v - a.side*sX;
    float * cb = f + (F_CBX+comp)*P;
    const float * eX = f + (F_EX+X)*P;
    const float decay = a.w0, drive = a.w1;
    if( isA ) {                              // cbY
      const float * eZ = f + (F_EX+Z)*P;
      float t1 = a.cdt_dX*( eZ[ vf - a.side*sX ] - eZ[vf] );
      t1 = a.side<0 ? t1 : -t1;
      float t2 = eX[ v + sZ - a.side*sX ];
      t2 = a.cdt_dZ*( t2 - eX[fh] );
      cb[v] = ( ( decay*cb[v] + drive*cb[fh] ) - t1 ) + t2;
    } else {                                 // cbZ
      const float * eY = f + (F_EX+Y)*P;
      float t1 = a.cdt_dX*( eY[ vf - a.side*sX ] - eY[vf] );
      t1 = a.side<0 ? t1 : -t1;
      float t2 = eX[ v + sY - a.side*sX ];
      t2 = a.cdt_dY*( t2 - eX[fh] );
      cb[v] = ( ( decay*cb[v] + drive*cb[fh] ) + t1 ) - t2;
    }
  } break;
  case OP_TANG_E_ZERO: f[ (F_EX+comp)*P + v ] = 0.f; f[ (F_TCAX+comp)*P + v ] = 0.f; break;
  case OP_JF_SCALE: { float * jf = f + (F_JFX+comp)*P; jf[v] = a.w0==0.f ? 0.f : jf[v]*a.w0; } break;
  case OP_NORM_B_ZERO: f[ (F_CBX+X)*P + v ] = 0.f; break;
  }
}

__global__ void __launch_bounds__(128)
k_plane_op( float * __restrict__ f, GridDev g, PlaneArgs a ) {
  plane_wait( a );
  plane_op_item( f, g, a, blockIdx.x*blockDim.x + threadIdx.x );
  plane_publish( a, a.done );
}

// Several INDEPENDENT plane operations in one launch (blockIdx.y picks the
// operation): the planes are a few hundred KB at most, so a step's ~24 single
// launches cost more in launch latency than in traffic.
constexpr int PLANE_BATCH = 12;
struct PlaneBatch { PlaneArgs a[PLANE_BATCH]; };

__global__ void __launch_bounds__(128)
k_plane_ops( float * __restrict__ f, GridDev g, PlaneBatch b ) {
  const PlaneArgs & a = b.a[blockIdx.y];
  plane_wait( a );
  plane_op_item( f, g, a, blockIdx.x*blockDim.x + threadIdx.x );
  plane_publish( a, a.done );
}

static int plane_items( const GridDev & g, int X, int op ) {
  const int n3[3] = { g.nx, g.ny, g.nz };
  const int nY = n3[(X+1)%3], nZ = n3[(X+2)%3];
  if( op==OP_NORM_B_ZERO ) return nY*nZ;
  return (nY+1)*nZ + nY*(nZ+1);
}

static void launch_plane( FieldMirror * m, PlaneArgs a ) {
  Runtime & r = rt();
  const GridDev & g = m->grid->dev;
  int n = plane_items( g, a.X, a.op );
  k_plane_op<<< div_up(n,128), 128, 0, r.stream >>>( m->d, g, a );
  VPB_CHECK_LAUNCH(); r.launches++;
}

// Collects operations that do not depend on one another and launches them together.
struct PlaneQueue {
  FieldMirror * m; PlaneBatch b; int n = 0, items = 0;
  explicit PlaneQueue( FieldMirror * m_ ) : m( m_ ) {}
  void add( const PlaneArgs & a ) {
    if( n==PLANE_BATCH ) flush();
    b.a[n++] = a;
    const int k = plane_items( m->grid->dev, a.X, a.op );
    if( k>items ) items = k;
  }
  void flush() {
    if( !n ) return;
    Runtime & r = rt();
    if( n==1 ) k_plane_op<<< div_up(items,128), 128, 0, r.stream >>>( m->d, m->grid->dev, b.a[0] );
    else       k_plane_ops<<< dim3( div_up(items,128), n ), 128, 0, r.stream >>>( m->d, m->grid->dev, b );
    VPB_CHECK_LAUNCH(); r.launches++;
    n = 0; items = 0;
  }
};

bool face_is_remote( const GridMirror * gm, int X, int side ) {
  int bc = gm->fbc[ X + ( side>0 ? 3 : 0 ) ];
  return bc>=0 && bc<g_world_size;
}
int face_bc( const GridMirror * gm, int X, int side ) { return gm->fbc[ X + ( side>0 ? 3 : 0 ) ]; }

// Face transport.  Single-GPU: what leaves through -X arrives on our own +X
// port (the reference self-sends through MPI, grid_comm.c:20-23), i.e. the
// receive buffer IS the opposite send buffer.  Multi-GPU: vpb_comm.cu.

// Device layout of m->d_face: [send_lo | send_hi | recv_lo | recv_hi], each face_floats long.
void exchange_axis( FieldMirror * m, int X, int nfloat, float ** from_lo, float ** from_hi, float * base ) {
  if( !base ) base = m->d_face;
  float * slo = base, * shi = base + m->face_floats;
  float * rlo = base + 2*m->face_floats, * rhi = base + 3*m->face_floats;
  const GridMirror * gm = m->grid;
  const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
  *from_lo = nullptr; *from_hi = nullptr;
  if( !rem_lo && !rem_hi ) return;
  const bool self_lo = rem_lo && face_bc( gm, X, -1 )==g_world_rank;
  const bool self_hi = rem_hi && face_bc( gm, X, +1 )==g_world_rank;
  if( ( rem_lo && !self_lo ) || ( rem_hi && !self_hi ) ) {
    comm_exchange_faces( m, X, rem_lo ? slo : nullptr, rem_hi ? shi : nullptr,
                         rem_lo ? rlo : nullptr, rem_hi ? rhi : nullptr, nfloat );
    if( rem_lo ) *from_lo = rlo;
    if( rem_hi ) *from_hi = rhi;
    return;
  }
  if( rem_lo ) *from_lo = shi;   // from the -X neighbour (ourselves): its +X message
  if( rem_hi ) *from_hi = slo;
}

// Fused variant of exchange_axis for remote (other-rank) neighbours: reserve the
// mailbox slots of this exchange so that the pack operations store straight into
// the neighbours' memory and the unpack operations read straight from ours.
// Returns false when the axis has no other-rank neighbour or the put transport
// cannot take the message (then the caller packs locally and calls exchange_axis).
static bool link_axis( FieldMirror * m, int X, int nfloat, CommLink * link ) {
  const GridMirror * gm = m->grid;
  const bool rem_lo = face_is_remote( gm, X, -1 ) && face_bc( gm, X, -1 )!=g_world_rank;
  const bool rem_hi = face_is_remote( gm, X, +1 ) && face_bc( gm, X, +1 )!=g_world_rank;
  if( !rem_lo && !rem_hi ) return false;
  // a face that wraps onto this very rank next to a face on another rank: leave it to exchange_axis
  if( ( face_is_remote( gm, X, -1 ) && !rem_lo ) || ( face_is_remote( gm, X, +1 ) && !rem_hi ) ) return false;
  return comm_link_reserve( rem_lo ? face_bc( gm, X, -1 ) : -1, rem_hi ? face_bc( gm, X, +1 ) : -1, sizeof(float)*(size_t)nfloat, link );
}
static void as_publisher( PlaneArgs & a, const CommLink & l, int side, int slot ) {
  a.buf = reinterpret_cast<float *>( l.dst[side] ); a.sync = 1; a.flag = l.pub_flag[side]; a.seq = l.pub_seq[side]; a.done = l.done + slot;
}
static void as_waiter( PlaneArgs & a, const CommLink & l, int side ) {
  a.buf = const_cast<float *>( reinterpret_cast<const float *>( l.src[side] ) ); a.sync = 2;
  a.flag = const_cast<unsigned *>( l.wait_flag[side] ); a.seq = l.wait_seq[side];
}

// remote.c:61-134 + local.c:50-122: fill every tangential-B ghost plane.
void fill_ghost_tang_b( FieldMirror * m ) {
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev; const vpb_grid_t * hg = gm->host;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const float d3[3] = { hg->dx, hg->dy, hg->dz };
  const float cdt[3] = { hg->cvac*hg->dt*hg->rdx, hg->cvac*hg->dt*hg->rdy, hg->cvac*hg->dt*hg->rdz };
  const float higend = ( g.nx>1 || g.ny>1 || g.nz>1 ) ? 1.03527618 : 1.;
  // Three phases instead of per-axis sequences: (1) every local ghost plane and
  // every message is written / packed in ONE launch -- the ghost planes of
  // different axes are disjoint, and the packs read face-normal and interior
  // values that no ghost operation writes; (2) the exchanges, each axis in its
  // own set of face buffers; (3) every received plane is unpacked in one launch.
  PlaneQueue q( m );
  CommLink link[3]; bool fused[3] = { false, false, false };
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    // local boundary conditions
    for( int side=-1; side<=1; side+=2 ) {
      const int bc = face_bc( gm, X, side );
      if( !( bc<0 || bc>=g_world_size ) ) continue;
      PlaneArgs a{}; a.X = X; a.side = side; a.plane = side<0 ? 0 : n3[X]+1;
      a.cdt_dX = cdt[X]; a.cdt_dY = cdt[Y]; a.cdt_dZ = cdt[Z];
      switch( bc ) {
      case VPB_PEC_FIELDS:       a.op = OP_GHOST_B_COPY; a.w0 =  1.f; break;
      case VPB_SYMMETRIC_FIELDS:
      case VPB_PMC_FIELDS:       a.op = OP_GHOST_B_COPY; a.w0 = -1.f; break;
      case VPB_ABSORB_FIELDS: {
        float drive = cdt[X]*higend;
        float decay = (1-drive)/(1+drive);
        drive = 2*drive/(1+drive);
        a.op = OP_GHOST_B_ABSORB; a.w0 = decay; a.w1 = drive;
      } break;
      default: VPB_ERROR( "Bad boundary condition encountered." );
      }
      q.add( a );
    }
    // remote (or periodic-to-self) neighbours; with the put transport the packs
    // store straight into the neighbours' mailboxes and publish (fused[X])
    float * base = m->d_face + (size_t)X*4*m->face_floats;
    fused[X] = link_axis( m, X, plane_items( g, X, OP_PACK_B ), &link[X] );
    if( face_is_remote( gm, X, -1 ) ) { PlaneArgs a{}; a.op = OP_PACK_B; a.X = X; a.side = -1; a.plane = 1;     a.buf = base;
                                        if( fused[X] && link[X].dst[0] ) as_publisher( a, link[X], 0, 2*X );   q.add( a ); }
    if( face_is_remote( gm, X, +1 ) ) { PlaneArgs a{}; a.op = OP_PACK_B; a.X = X; a.side = +1; a.plane = n3[X]; a.buf = base + m->face_floats;
                                        if( fused[X] && link[X].dst[1] ) as_publisher( a, link[X], 1, 2*X+1 ); q.add( a ); }
  }
  q.flush();
  float * from[3][2] = {};
  for( int X=0; X<3; X++ ) {
    if( fused[X] || ( !face_is_remote( gm, X, -1 ) && !face_is_remote( gm, X, +1 ) ) ) continue;
    exchange_axis( m, X, plane_items( g, X, OP_PACK_B ), &from[X][0], &from[X][1], m->d_face + (size_t)X*4*m->face_floats );
  }
  for( int X=0; X<3; X++ ) {
    // equal cell size on both sides (checked when the communicator is set up)
    float lw = d3[X];
    const float rw = (2.*d3[X])/(lw+d3[X]);
    lw = (lw-d3[X])/(lw+d3[X]);
    for( int sd=0; sd<2; sd++ ) {
      const bool waits = fused[X] && link[X].src[sd];
      if( !waits && !from[X][sd] ) continue;
      PlaneArgs a{}; a.op = OP_UNPACK_GHOST_B; a.X = X; a.side = sd ? +1 : -1; a.plane = sd ? n3[X]+1 : 0; a.w0 = rw; a.w1 = lw; a.buf = from[X][sd];
      if( waits ) as_waiter( a, link[X], sd );
      q.add( a );
    }
  }
  q.flush();
}

// local.c:267-296
void local_adjust_norm_b( FieldMirror * m ) {
  const GridDev & g = m->grid->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( m->grid, X, side );
    if( !( bc<0 || bc>=g_world_size ) ) continue;
    if( bc==VPB_SYMMETRIC_FIELDS ) {
      PlaneArgs a{}; a.op = OP_NORM_B_ZERO; a.X = X; a.side = side; a.plane = side<0 ? 1 : n3[X]+1;
      launch_plane( m, a );
    } else if( bc!=VPB_PEC_FIELDS && bc!=VPB_PMC_FIELDS && bc!=VPB_ABSORB_FIELDS )
      VPB_ERROR( "Bad boundary condition encountered." );
  }
}

// local.c:224-264
void local_adjust_tang_e( FieldMirror * m ) {
  const GridDev & g = m->grid->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( m->grid, X, side );
    if( !( bc<0 || bc>=g_world_size ) ) continue;
    if( bc==VPB_PEC_FIELDS ) {
      PlaneArgs a{}; a.op = OP_TANG_E_ZERO; a.X = X; a.side = side; a.plane = side<0 ? 1 : n3[X]+1;
      launch_plane( m, a );
    }
  }
}

} // namespace vpb

using namespace vpb;


extern "C" {

// advance_b.cc:10-19 -> advance_b_pipeline.cc:52-125 (+ local_adjust_norm_b)
void vpic_b200_advance_b( vpb_field_array_t * fa, float frac ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * m = mirror( fa );
  to_device( m );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const float px = (g.nx>1) ? frac*hg->cvac*hg->dt*hg->rdx : 0;
  const float py = (g.ny>1) ? frac*hg->cvac*hg->dt*hg->rdy : 0;
  const float pz = (g.nz>1) ? frac*hg->cvac*hg->dt*hg->rdz : 0;
  k_advance_b<<< dim3( div_up( (long long)(g.nx+1)*(g.ny+1), 256 ), g.nz+1 ), 256, 0, r.stream >>>( m->d, g, px, py, pz );
  VPB_CHECK_LAUNCH(); r.launches++;
  local_adjust_norm_b( m );
  finish_write( m );
}

// advance_e.cc:10-25 / vacuum_advance_e.cc:10-25
void vpic_b200_advance_e( vpb_field_array_t * fa, float frac ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  if( frac!=1 ) VPB_ERROR( "standard advance_e does not support frac != 1 yet" );
  Runtime & r = rt();
  FieldMirror * m = mirror( fa );
  to_device( m );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const float damp = prm->damp;
  fill_ghost_tang_b( m );
  if( m->n_mc==1 ) {
    const vpb_material_coefficient_t * mc = prm->mc;
    VacuumCoef c;
    c.decayx = mc->decayx; c.drivex = mc->drivex; c.decayy = mc->decayy; c.drivey = mc->drivey;
    c.decayz = mc->decayz; c.drivez = mc->drivez;
    c.px_muz = ((g.nx>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdx : 0)*mc->rmuz;
    c.px_muy = ((g.nx>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdx : 0)*mc->rmuy;
    c.py_mux = ((g.ny>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdy : 0)*mc->rmux;
    c.py_muz = ((g.ny>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdy : 0)*mc->rmuz;
    c.pz_muy = ((g.nz>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdz : 0)*mc->rmuy;
    c.pz_mux = ((g.nz>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdz : 0)*mc->rmux;
    c.damp = damp; c.cj = hg->dt/hg->eps0;
    k_advance_e_vacuum<<< dim3( div_up( (long long)(g.nx+1)*(g.ny+1), 256 ), g.nz+1 ), 256, 0, r.stream >>>( m->d, g, c );
  } else {
    const float px = (g.nx>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdx : 0;
    const float py = (g.ny>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdy : 0;
    const float pz = (g.nz>1) ? (1+damp)*hg->cvac*hg->dt*hg->rdz : 0;
    k_advance_e_material<<< dim3( div_up( (long long)(g.nx+1)*(g.ny+1), 256 ), g.nz+1 ), 256, 0, r.stream >>>( m->d, m->d_mat, m->d_mc, g, px, py, pz, damp, hg->dt/hg->eps0 );
  }
  VPB_CHECK_LAUNCH(); r.launches++;
  local_adjust_tang_e( m );
  finish_write( m );
}

// energy_f.c:12-23 / vacuum_energy_f.c
void vpic_b200_energy_f( double * en, const vpb_field_array_t * fa ) {
  if( !en || !fa ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * m = mirror( const_cast<vpb_field_array_t *>( fa ) );
  to_device( m );
  const GridDev & g = m->grid->dev; const vpb_grid_t * hg = fa->g;
  if( r.deposit_mode==VPB_DEPOSIT_PARITY ) {      // the reference's summation order, bit for bit
    double loc[6];
    energy_f_parity( loc, m );
    if( g_world_size>1 ) comm_allsum_d( loc, 6 );
    for( int c=0; c<6; c++ ) en[c] = loc[c];
    return;
  }
  const int ncell = g.nx*g.ny*g.nz;
  int nb = div_up( ncell, EN_BLOCK ); if( nb>r.sm_count*4 ) nb = r.sm_count*4;
  double * partial = (double *)scratch( sizeof(double)*6*(size_t)nb );
  if( m->n_mc==1 ) k_energy_f<true ><<< nb, EN_BLOCK, 0, r.stream >>>( partial, m->d, m->d_mat, m->d_mc, g );
  else             k_energy_f<false><<< nb, EN_BLOCK, 0, r.stream >>>( partial, m->d, m->d_mat, m->d_mc, g );
  VPB_CHECK_LAUNCH(); r.launches++;
  const double v0 = 0.5*hg->eps0*hg->dV;
  k_sum_partials<<< 1, 32, 0, r.stream >>>( m->d_en, partial, nb, 6, v0 );
  VPB_CHECK_LAUNCH(); r.launches++;
  double local[6];
  d2h( local, m->d_en, sizeof(double)*6, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( g_world_size>1 ) comm_allsum_d( local, 6 );
  for( int c=0; c<6; c++ ) en[c] = local[c];
}

// sfa.c:231-237 (the jfx,jfy,jfz planes are adjacent in the SoA block)
void vpic_b200_clear_jf( vpb_field_array_t * fa ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * m = mirror( fa );
  to_device( m );
  VPB_CUDA( cudaMemsetAsync( m->comp( F_JFX ), 0, sizeof(float)*3*(size_t)m->grid->dev.nvp, r.stream ) );
  r.launches++;
  finish_write( m );
}

void vpic_b200_clear_rhof( vpb_field_array_t * fa ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * m = mirror( fa );
  to_device( m );
  VPB_CUDA( cudaMemsetAsync( m->comp( F_RHOF ), 0, sizeof(float)*(size_t)m->grid->dev.nvp, r.stream ) );
  r.launches++;
  finish_write( m );
}

// remote.c:417-508 (+ local_adjust_jf, local.c:335-368)
void vpic_b200_synchronize_jf( vpb_field_array_t * fa ) {
  if( !fa ) VPB_ERROR( "Bad args" );
  FieldMirror * m = mirror( fa );
  to_device( m );
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev; const vpb_grid_t * hg = fa->g;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const float d3[3] = { hg->dx, hg->dy, hg->dz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    const int bc = face_bc( gm, X, side );
    if( !( bc<0 || bc>=g_world_size ) ) continue;
    PlaneArgs a{}; a.op = OP_JF_SCALE; a.X = X; a.side = side; a.plane = side<0 ? 1 : n3[X]+1;
    if( bc==VPB_PEC_FIELDS ) a.w0 = 0.f;
    else if( bc==VPB_SYMMETRIC_FIELDS || bc==VPB_PMC_FIELDS || bc==VPB_ABSORB_FIELDS ) a.w0 = 2.f;
    else VPB_ERROR( "Bad boundary condition encountered." );
    launch_plane( m, a );
  }
  for( int X=0; X<3; X++ ) {
    const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
    if( !rem_lo && !rem_hi ) continue;
    const int nf = plane_items( g, X, OP_PACK_J );
    // both messages are packed before either is folded in (remote.c:474-481)
    // (the two packs read, and the two folds write, different planes: one launch each;
    // the axes stay in sequence because edges shared by two faces are folded twice)
    PlaneQueue q( m );
    CommLink link; const bool fused = link_axis( m, X, nf, &link );
    if( rem_lo ) { PlaneArgs a{}; a.op = OP_PACK_J; a.X = X; a.side = -1; a.plane = 1;       a.buf = m->d_face;
                   if( fused && link.dst[0] ) as_publisher( a, link, 0, 0 ); q.add( a ); }
    if( rem_hi ) { PlaneArgs a{}; a.op = OP_PACK_J; a.X = X; a.side = +1; a.plane = n3[X]+1; a.buf = m->d_face + m->face_floats;
                   if( fused && link.dst[1] ) as_publisher( a, link, 1, 1 ); q.add( a ); }
    q.flush();
    float * from_lo = nullptr, * from_hi = nullptr;
    if( !fused ) exchange_axis( m, X, nf, &from_lo, &from_hi );
    float rw = d3[X], lw = rw + d3[X];
    rw /= lw; lw = d3[X]/lw; lw += lw; rw += rw;
    // from the -X neighbour (its plane nX+1) -> our plane 1; from +X (its plane 1) -> our plane nX+1
    if( from_lo || ( fused && link.src[0] ) ) { PlaneArgs a{}; a.op = OP_FOLD_J; a.X = X; a.side = -1; a.plane = 1;       a.w0 = rw; a.w1 = lw; a.buf = from_lo;
                                                 if( fused ) as_waiter( a, link, 0 ); q.add( a ); }
    if( from_hi || ( fused && link.src[1] ) ) { PlaneArgs a{}; a.op = OP_FOLD_J; a.X = X; a.side = +1; a.plane = n3[X]+1; a.w0 = rw; a.w1 = lw; a.buf = from_hi;
                                                 if( fused ) as_waiter( a, link, 1 ); q.add( a ); }
    q.flush();
  }
  finish_write( m );
}

} // extern "C"
