// vpb_boundary.cu -- boundary_p: particles that left the rank's domain.
// Reference: src/boundary/boundary_p.cc:36-518 and accumulate_rhob,
// src/species_advance/standard/rho_p.cc:125-171.
//
// The reference walks each species' movers in REVERSE order, turns every one
// into "absorbed" (charge to rhob) or "sent" (a 48 B injector appended to the
// face's message), back-fills the hole with the last particle, then appends
// the injectors it received (faces 0..5, each message in reverse) and finishes
// their displacement with move_p.  That sequential recipe fixes the order of
// the particle array afterwards, which sort_p's stable output depends on, so
// the device version reproduces it exactly, in parallel:
//   * the k-th mover in processing order gets the k-th slot of its face's
//     message (block-wide ranking per chunk of 1024 movers),
//   * all movers leave, so every surviving particle of the tail follows its own
//     chain of "p[i] = p[--np]" copies to a hole below the new end of the array
//     (k_bnd_backfill),
//   * received injectors are ranked per species the same way and moved with
//     every lane busy.
// Messages travel as device buffers over NCCL (vpb_comm.cu): counts, then
// payloads, like boundary_p.cc:292-316.
#include "vpb_push.cuh"
#include <vector>

namespace vpb {

constexpr int BND_T = 1024;           // threads of the single ranking block
constexpr int BND_MAX_SP = 16;        // species ids the inject kernel can address

struct BndFaces {
  int kind[6];                        // ParticleFaceKind
  int code[6];                        // regular box: the face's neighbor[] code (handler id) if kind is PK_CUSTOM
  unsigned handler_mask;              // bit h set: the host program registered the handler with id -h-3
  int nb_n[6];                        // neighbour extent along the face axis
  vpb_particle_injector_t * send[6];  // device message bodies
  float4 * custom;                    // particles that hit a face with a user handler: 4 x float4 each
  int * face_count;                   // [7] injectors packed so far (all species); [6] = custom entries
};

// Ranks of the threads of a block among those holding the same key (0..NKEY-1,
// or <0 for "none"), in thread order.  Returns the rank and leaves the
// per-key totals of this chunk in s_tot.
template<int NKEY>
__device__ __forceinline__ int block_rank_by_key( int key, int (*s_w)[NKEY], int * s_tot ) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for( int t=threadIdx.x; t<(BND_T/32)*NKEY; t+=BND_T ) (&s_w[0][0])[t] = 0;
  __syncthreads();
  const unsigned m = __match_any_sync( 0xffffffffu, key );
  const int r_in_warp = __popc( m & ( ( 1u<<lane )-1u ) );
  if( key>=0 && r_in_warp==0 ) s_w[w][key] = __popc( m );
  __syncthreads();
  int before = 0;
  if( key>=0 ) for( int k=0; k<w; k++ ) before += s_w[k][key];
  if( threadIdx.x<NKEY ) { int t = 0; for( int k=0; k<BND_T/32; k++ ) t += s_w[k][threadIdx.x]; s_tot[threadIdx.x] = t; }
  __syncthreads();
  return before + r_in_warp;
}

// accumulate_rhob, rho_p.cc:125-171 (separate multiply and add, as compiled).
__device__ __forceinline__ void accumulate_rhob_dev( float * __restrict__ rhob, const GridDev & g,
                                                     float dx, float dy, float dz, int v, float w, float q_r8V ) {
  float w0 = dx, w1 = dy, w2, w3, w4, w5, w6, w7 = q_r8V*w;
  w6 = w7 - w0*w7; w7 = w7 + w0*w7;
  w4 = w6 - w1*w6; w5 = w7 - w1*w7;
  w6 = w6 + w1*w6; w7 = w7 + w1*w7;
  w0 = w4 - dz*w4; w1 = w5 - dz*w5; w2 = w6 - dz*w6; w3 = w7 - dz*w7;
  w4 = w4 + dz*w4; w5 = w5 + dz*w5; w6 = w6 + dz*w6; w7 = w7 + dz*w7;
  int x = v; const int z = x/g.sz;
  if( z==1    ) { w0 += w0; w1 += w1; w2 += w2; w3 += w3; }
  if( z==g.nz ) { w4 += w4; w5 += w5; w6 += w6; w7 += w7; }
  x -= g.sz*z; const int y = x/g.sy;
  if( y==1    ) { w0 += w0; w1 += w1; w4 += w4; w5 += w5; }
  if( y==g.ny ) { w2 += w2; w3 += w3; w6 += w6; w7 += w7; }
  x -= g.sy*y;
  if( x==1    ) { w0 += w0; w2 += w2; w4 += w4; w6 += w6; }
  if( x==g.nx ) { w1 += w1; w3 += w3; w5 += w5; w7 += w7; }
  atomicAdd( rhob + v,           w0 ); atomicAdd( rhob + v+1,           w1 );
  atomicAdd( rhob + v+g.sy,      w2 ); atomicAdd( rhob + v+g.sy+1,      w3 );
  atomicAdd( rhob + v+g.sz,      w4 ); atomicAdd( rhob + v+g.sz+1,      w5 );
  atomicAdd( rhob + v+g.sz+g.sy, w6 ); atomicAdd( rhob + v+g.sz+g.sy+1, w7 );
}

// One species: absorb or pack every mover, in the reference's processing
// order (last mover first).  Single block; chunks of BND_T movers.
__global__ void __launch_bounds__(BND_T)
k_bnd_classify( const vpb_particle_t * __restrict__ p, const vpb_particle_mover_t * __restrict__ pm, int nm,
                int sp_id, float q_r8V, float * __restrict__ rhob, GridDev g, BndFaces a, int * __restrict__ n_unknown,
                bool ordered ) {
  __shared__ int s_w[BND_T/32][7];
  __shared__ int s_tot[7], s_base[7];
  if( threadIdx.x<7 ) s_base[threadIdx.x] = a.face_count[threadIdx.x];
  __syncthreads();
  for( int c0=0; c0<nm; c0+=BND_T ) {
    const int r = c0 + threadIdx.x;              // position in processing order
    int key = -1, face = 0, voxel = 0, i = 0, code = 0;
    bool absorbed = false;
    float4 r0 = {}, r1 = {}, d = {};
    if( r<nm ) {
      d = reinterpret_cast<const float4 *>( pm )[ nm-1-r ];
      i = __float_as_int( d.w );
      r0 = reinterpret_cast<const float4 *>( p + i )[0];
      r1 = reinterpret_cast<const float4 *>( p + i )[1];
      const int vi = __float_as_int( r0.w );
      face = vi & 7; voxel = vi >> 3;
      // what the host table says about this exit (boundary_p.cc:196-260): the
      // face's kind on a regular box, the voxel's own code otherwise
      int kind = -1;
      if( face<6 ) {
        if( g.neighbor ) {
          code = g.neighbor[ 6*(size_t)voxel + face ];
          kind = code==NB_ABSORB ? PK_ABSORB : code==NB_REMOTE ? PK_REMOTE : ( code<=-3 && code>NB_UNKNOWN ) ? PK_CUSTOM : -1;
        } else { kind = a.kind[face]; code = a.code[face]; }
      }
      if( kind==PK_ABSORB ) { if( ordered ) absorbed = true; else accumulate_rhob_dev( rhob, g, r0.x, r0.y, r0.z, voxel, r1.w, q_r8V ); }
      else if( kind==PK_REMOTE ) key = face;
      else if( kind==PK_CUSTOM && a.custom && code>=-34 && ( ( a.handler_mask >> ( -code-3 ) ) & 1u ) ) key = 6;
      else atomicAdd( n_unknown, 1 );
    }
    if( ordered ) {
      // PARITY: the absorbed charge lands on rhob one particle after the other
      // in processing order, like the reference's loop (boundary_p.cc:204-208)
      for( int w=0; w<BND_T/32; w++ ) {
        if( ( threadIdx.x>>5 )==w ) {
          unsigned todo = __ballot_sync( 0xffffffffu, absorbed );
          while( todo ) {
            const int l = __ffs( todo )-1; todo &= todo-1;
            if( ( threadIdx.x&31 )==l ) accumulate_rhob_dev( rhob, g, r0.x, r0.y, r0.z, voxel, r1.w, q_r8V );
            __syncwarp();
          }
        }
        __syncthreads();
      }
    }
    const int rank = block_rank_by_key<7>( key, s_w, s_tot );
    if( key==6 ) {
      // handed to the host program's handler in processing order (boundary_p.cc:236-253);
      // the handler sees the particle with its voxel restored (:196-199)
      float4 * out = a.custom + 4*(size_t)( s_base[6] + rank );
      out[0] = make_float4( r0.x, r0.y, r0.z, __int_as_float( voxel ) );
      out[1] = r1;
      out[2] = d;
      out[3] = make_float4( __int_as_float( face ), __int_as_float( sp_id ), __int_as_float( code ), 0.f );
    } else if( key>=0 ) {
      // the same voxel seen from the neighbour brick (boundary_p.cc:215-230)
      const int axis = face>=3 ? face-3 : face;
      int c[3] = { voxel % g.sy, ( voxel / g.sy ) % ( g.ny+2 ), voxel / g.sz };
      int rn[3] = { g.nx, g.ny, g.nz };
      rn[axis] = a.nb_n[face];
      c[axis] = face<3 ? rn[axis] : 1;
      const int nv_i = c[0] + ( rn[0]+2 )*( c[1] + ( rn[1]+2 )*c[2] );
      float pos[3] = { r0.x, r0.y, r0.z };
      pos[axis] = face<3 ? 1.f : -1.f;
      float4 * out = reinterpret_cast<float4 *>( a.send[face] + s_base[face] + rank );
      out[0] = make_float4( pos[0], pos[1], pos[2], __int_as_float( nv_i ) );
      out[1] = r1;
      out[2] = make_float4( d.x, d.y, d.z, __int_as_float( sp_id ) );
    }
    __syncthreads();
    if( threadIdx.x<7 ) s_base[threadIdx.x] += s_tot[threadIdx.x];
    __syncthreads();
  }
  if( threadIdx.x<7 ) a.face_count[threadIdx.x] = s_base[threadIdx.x];
}

// All nm movers (ascending particle index in pm) leave.  The reference's
// "p[i] = p[--np]" walked over the movers in reverse means: step k (k=1..nm)
// copies position np-k onto the k-th largest leaving index r_k.  A surviving
// particle of the tail [new_np, np) at position np-k therefore lands on r_k;
// if r_k is itself inside the tail it is picked up again at step np-r_k, and so
// on until it reaches a hole below new_np.  Every tail survivor follows its
// own chain, independently of the others.
__device__ __forceinline__ int lower_bound_idx( const vpb_particle_mover_t * pm, int n, int key ) {   // first m with pm[m].i >= key
  int lo = 0, hi = n;
  while( lo<hi ) { const int mid = ( lo+hi )>>1; if( pm[mid].i<key ) lo = mid+1; else hi = mid; }
  return lo;
}

__global__ void __launch_bounds__(256)
k_bnd_backfill( vpb_particle_t * __restrict__ p, const vpb_particle_mover_t * __restrict__ pm, int nm, int np ) {
  const int new_np = np - nm;
  const int j = new_np + blockIdx.x*blockDim.x + threadIdx.x;
  if( j>=np ) return;
  const int lb = lower_bound_idx( pm, nm, j );
  if( lb<nm && pm[lb].i==j ) return;                 // j itself leaves
  int k = np - j, hole;
  for(;;) {
    hole = pm[ nm-k ].i;                             // r_k
    if( hole<new_np ) break;
    k = np - hole;
  }
  const float4 a = reinterpret_cast<const float4 *>( p + j )[0], b = reinterpret_cast<const float4 *>( p + j )[1];
  reinterpret_cast<float4 *>( p + hole )[0] = a;
  reinterpret_cast<float4 *>( p + hole )[1] = b;
}

// advance_p wrote the voxel key of every particle for the lazy sort_p (vpb_sort.cu) before boundary_p moved things:
// refresh the keys of the records boundary_p wrote -- the holes a tail particle filled (movers, n>0: the indices
// pm[k].i below new_np) or the records appended by the injection (pm==nullptr: the range [first,first+n)).
__global__ void __launch_bounds__(256)
k_keys_patch( int * __restrict__ keys, const vpb_particle_t * __restrict__ p, const vpb_particle_mover_t * __restrict__ pm, int n, int first, int new_np ) {
  const int k = blockIdx.x*blockDim.x + threadIdx.x;
  if( k>=n ) return;
  const int i = pm ? pm[k].i : first + k;
  if( i<new_np ) keys[i] = p[i].i;
}

struct BndInject {
  const vpb_particle_injector_t * recv[7];
  int start[8];                        // prefix sums of the counts: local handler injectors, then faces 0..5
  vpb_particle_t * p[BND_MAX_SP];
  vpb_particle_mover_t * pm[BND_MAX_SP];
  int np[BND_MAX_SP], nm[BND_MAX_SP], max_nm[BND_MAX_SP];
  float q[BND_MAX_SP];
  int * counters;                      // [0..MAX_SP) injected, [MAX_SP..2 MAX_SP) movers appended, [2 MAX_SP..) movers dropped
};

// One injected particle: appended at idx, its remaining displacement finished
// with move_p (currents into the accumulator, a new mover if it leaves again).
__device__ __forceinline__ void inject_one( const BndInject & a, float * __restrict__ acc, const GridDev & g,
                                            float4 i0, float4 i1, float4 i2, int key, int idx ) {
  MoveState ms;
  ms.dx = i0.x; ms.dy = i0.y; ms.dz = i0.z; ms.i = __float_as_int( i0.w );
  ms.ux = i1.x; ms.uy = i1.y; ms.uz = i1.z;
  ms.dispx = i2.x; ms.dispy = i2.y; ms.dispz = i2.z;
  const float q = a.q[key]*i1.w;
  float j[12]; int status;
  do {
    int vox;
    status = move_p_streak( ms, q, g, j, vox );
    commit_currents( acc, vox, j );
  } while( status==1 );
  float4 * dst = reinterpret_cast<float4 *>( a.p[key] + idx );
  dst[0] = make_float4( ms.dx, ms.dy, ms.dz, __int_as_float( ms.i ) );
  dst[1] = make_float4( ms.ux, ms.uy, ms.uz, i1.w );
  if( status==2 ) {
    const int k = a.nm[key] + atomicAdd( a.counters + BND_MAX_SP + key, 1 );
    if( k<a.max_nm[key] ) reinterpret_cast<float4 *>( a.pm[key] )[k] = make_float4( ms.dispx, ms.dispy, ms.dispz, __int_as_float( idx ) );
    else atomicAdd( a.counters + 2*BND_MAX_SP + key, 1 );
  }
}

// PARITY deposit: the same injection by one thread, so that the currents of
// injected particles are added to the accumulator in the reference's order.
__global__ void k_bnd_inject_serial( BndInject a, float * __restrict__ acc, GridDev g ) {
  if( threadIdx.x || blockIdx.x ) return;
  int count[BND_MAX_SP];
  for( int k=0; k<BND_MAX_SP; k++ ) count[k] = 0;
  const int total = a.start[7];
  int face = 0;
  for( int r=0; r<total; r++ ) {
    while( r>=a.start[face+1] ) face++;
    const int n = a.start[face+1] - a.start[face];
    const float4 * src = reinterpret_cast<const float4 *>( a.recv[face] + ( n-1-( r-a.start[face] ) ) );
    const float4 i0 = src[0], i1 = src[1], i2 = src[2];
    const int key = __float_as_int( i2.w );
    if( key<0 || key>=BND_MAX_SP || !a.p[key] ) { atomicAdd( a.counters + 3*BND_MAX_SP, 1 ); continue; }
    inject_one( a, acc, g, i0, i1, i2, key, a.np[key] + count[key] );
    count[key]++;
  }
  for( int k=0; k<BND_MAX_SP; k++ ) a.counters[k] = count[k];
}

// Append the injected particles (the local handlers' injectors first, then the
// messages of faces 0..5, each list last-first) to their species and finish
// their move (boundary_p.cc:426-494).
__global__ void __launch_bounds__(BND_T)
k_bnd_inject( BndInject a, float * __restrict__ acc, GridDev g ) {
  __shared__ int s_w[BND_T/32][BND_MAX_SP];
  __shared__ int s_tot[BND_MAX_SP], s_base[BND_MAX_SP];
  if( threadIdx.x<BND_MAX_SP ) s_base[threadIdx.x] = 0;
  __syncthreads();
  const int total = a.start[7];
  for( int c0=0; c0<total; c0+=BND_T ) {
    const int r = c0 + threadIdx.x;
    int key = -1; float4 i0 = {}, i1 = {}, i2 = {};
    if( r<total ) {
      int face = 0;
      while( r>=a.start[face+1] ) face++;
      const int n = a.start[face+1] - a.start[face];
      const float4 * src = reinterpret_cast<const float4 *>( a.recv[face] + ( n-1-( r-a.start[face] ) ) );
      i0 = src[0]; i1 = src[1]; i2 = src[2];
      key = __float_as_int( i2.w );
      if( key<0 || key>=BND_MAX_SP || !a.p[key] ) key = -1;      // unknown species: dropped (counted below)
    }
    const int rank = block_rank_by_key<BND_MAX_SP>( key, s_w, s_tot );
    if( key>=0 ) inject_one( a, acc, g, i0, i1, i2, key, a.np[key] + s_base[key] + rank );
    else if( r<total ) atomicAdd( a.counters + 3*BND_MAX_SP, 1 );
    __syncthreads();
    if( threadIdx.x<BND_MAX_SP ) s_base[threadIdx.x] += s_tot[threadIdx.x];
    __syncthreads();
  }
  if( threadIdx.x<BND_MAX_SP ) a.counters[threadIdx.x] = s_base[threadIdx.x];
}

// ---- multi-block versions (FAST deposit) ------------------------------------------
// The single-block kernels above walk the movers 1024 at a time, one chunk after
// the other; at config 4's face sizes (10^5 movers per species and step) that
// serial walk was the largest part of boundary_p.  Here every chunk of BND_T
// movers is a block of its own: the block ranks its movers per key, then takes
// the running per-key totals from the chunk before it and publishes its own
// (a chained scan: chunk numbers are handed out by an atomic ticket, so the
// chunk a block waits for has always started).  Slots are the same as the
// single-block kernel's, so message and array order stay bit-exact.
struct ChainScan {
  int * ticket;          // [1]
  int * flag;            // [nchunk]  1 = prefix[c] is published
  int * prefix;          // [nchunk][16] inclusive per-key totals
};

template<int NKEY>
__device__ __forceinline__ void chain_bases( const ChainScan & cs, int chunk, int nchunk, const int * s_tot, int * s_base,
                                             const int * first_base, int * last_out ) {
  if( threadIdx.x<NKEY ) {
    int base;
    if( chunk==0 ) base = first_base ? first_base[threadIdx.x] : 0;
    else {
      const volatile int * fl = cs.flag + ( chunk-1 );
      while( *fl==0 ) __nanosleep( 20 );
      __threadfence();
      base = reinterpret_cast<const volatile int *>( cs.prefix )[ 16*(size_t)( chunk-1 ) + threadIdx.x ];
    }
    s_base[threadIdx.x] = base;
    const int incl = base + s_tot[threadIdx.x];
    cs.prefix[ 16*(size_t)chunk + threadIdx.x ] = incl;
    if( chunk==nchunk-1 && last_out ) last_out[threadIdx.x] = incl;
    __threadfence();
  }
  __syncthreads();
  if( threadIdx.x==0 ) { __threadfence(); reinterpret_cast<volatile int *>( cs.flag )[chunk] = 1; }
}

__global__ void __launch_bounds__(BND_T)
k_bnd_classify_mb( const vpb_particle_t * __restrict__ p, const vpb_particle_mover_t * __restrict__ pm, int nm,
                   int sp_id, float q_r8V, float * __restrict__ rhob, GridDev g, BndFaces a, int * __restrict__ n_unknown,
                   ChainScan cs, int nchunk ) {
  __shared__ int s_w[BND_T/32][7];
  __shared__ int s_tot[7], s_base[7], s_chunk;
  if( threadIdx.x==0 ) s_chunk = atomicAdd( cs.ticket, 1 );
  __syncthreads();
  const int chunk = s_chunk;
  const int r = chunk*BND_T + threadIdx.x;       // position in processing order
  int key = -1, face = 0, voxel = 0, i = 0, code = 0;
  float4 r0 = {}, r1 = {}, d = {};
  if( r<nm ) {
    d = reinterpret_cast<const float4 *>( pm )[ nm-1-r ];
    i = __float_as_int( d.w );
    r0 = reinterpret_cast<const float4 *>( p + i )[0];
    r1 = reinterpret_cast<const float4 *>( p + i )[1];
    const int vi = __float_as_int( r0.w );
    face = vi & 7; voxel = vi >> 3;
    int kind = -1;
    if( face<6 ) {
      if( g.neighbor ) {
        code = g.neighbor[ 6*(size_t)voxel + face ];
        kind = code==NB_ABSORB ? PK_ABSORB : code==NB_REMOTE ? PK_REMOTE : ( code<=-3 && code>NB_UNKNOWN ) ? PK_CUSTOM : -1;
      } else { kind = a.kind[face]; code = a.code[face]; }
    }
    if( kind==PK_ABSORB ) accumulate_rhob_dev( rhob, g, r0.x, r0.y, r0.z, voxel, r1.w, q_r8V );
    else if( kind==PK_REMOTE ) key = face;
    else if( kind==PK_CUSTOM && a.custom && code>=-34 && ( ( a.handler_mask >> ( -code-3 ) ) & 1u ) ) key = 6;
    else atomicAdd( n_unknown, 1 );
  }
  const int rank = block_rank_by_key<7>( key, s_w, s_tot );
  chain_bases<7>( cs, chunk, nchunk, s_tot, s_base, a.face_count, a.face_count );
  if( key==6 ) {
    float4 * out = a.custom + 4*(size_t)( s_base[6] + rank );
    out[0] = make_float4( r0.x, r0.y, r0.z, __int_as_float( voxel ) );
    out[1] = r1;
    out[2] = d;
    out[3] = make_float4( __int_as_float( face ), __int_as_float( sp_id ), __int_as_float( code ), 0.f );
  } else if( key>=0 ) {
    const int axis = face>=3 ? face-3 : face;
    int c[3] = { voxel % g.sy, ( voxel / g.sy ) % ( g.ny+2 ), voxel / g.sz };
    int rn[3] = { g.nx, g.ny, g.nz };
    rn[axis] = a.nb_n[face];
    c[axis] = face<3 ? rn[axis] : 1;
    const int nv_i = c[0] + ( rn[0]+2 )*( c[1] + ( rn[1]+2 )*c[2] );
    float pos[3] = { r0.x, r0.y, r0.z };
    pos[axis] = face<3 ? 1.f : -1.f;
    float4 * out = reinterpret_cast<float4 *>( a.send[face] + s_base[face] + rank );
    out[0] = make_float4( pos[0], pos[1], pos[2], __int_as_float( nv_i ) );
    out[1] = r1;
    out[2] = make_float4( d.x, d.y, d.z, __int_as_float( sp_id ) );
  }
}

__global__ void __launch_bounds__(BND_T)
k_bnd_inject_mb( BndInject a, float * __restrict__ acc, GridDev g, ChainScan cs, int nchunk ) {
  __shared__ int s_w[BND_T/32][BND_MAX_SP];
  __shared__ int s_tot[BND_MAX_SP], s_base[BND_MAX_SP], s_chunk;
  if( threadIdx.x==0 ) s_chunk = atomicAdd( cs.ticket, 1 );
  __syncthreads();
  const int chunk = s_chunk;
  const int total = a.start[7];
  const int r = chunk*BND_T + threadIdx.x;
  int key = -1; float4 i0 = {}, i1 = {}, i2 = {};
  if( r<total ) {
    int face = 0;
    while( r>=a.start[face+1] ) face++;
    const int n = a.start[face+1] - a.start[face];
    const float4 * src = reinterpret_cast<const float4 *>( a.recv[face] + ( n-1-( r-a.start[face] ) ) );
    i0 = src[0]; i1 = src[1]; i2 = src[2];
    key = __float_as_int( i2.w );
    if( key<0 || key>=BND_MAX_SP || !a.p[key] ) key = -1;
  }
  const int rank = block_rank_by_key<BND_MAX_SP>( key, s_w, s_tot );
  chain_bases<BND_MAX_SP>( cs, chunk, nchunk, s_tot, s_base, nullptr, a.counters );
  if( key>=0 ) inject_one( a, acc, g, i0, i1, i2, key, a.np[key] + s_base[key] + rank );
  else if( r<total ) atomicAdd( a.counters + 3*BND_MAX_SP, 1 );
}

// STRICT coherence, sparse transfers: rows of the particle array named by an
// index list, device <-> a compact staging buffer.
__global__ void __launch_bounds__(256)
k_rows_scatter( vpb_particle_t * __restrict__ p, const int * __restrict__ idx, const vpb_particle_t * __restrict__ rows, int n ) {
  const int k = blockIdx.x*blockDim.x + threadIdx.x;
  if( k>=n ) return;
  const float4 * s = reinterpret_cast<const float4 *>( rows + k );
  float4 * d = reinterpret_cast<float4 *>( p + idx[k] );
  d[0] = s[0]; d[1] = s[1];
}
__global__ void __launch_bounds__(256)
k_rows_gather( vpb_particle_t * __restrict__ rows, const int * __restrict__ idx, const vpb_particle_t * __restrict__ p, int n ) {
  const int k = blockIdx.x*blockDim.x + threadIdx.x;
  if( k>=n ) return;
  const float4 * s = reinterpret_cast<const float4 *>( p + idx[k] );
  float4 * d = reinterpret_cast<float4 *>( rows + k );
  d[0] = s[0]; d[1] = s[1];
}

} // namespace vpb

using namespace vpb;

namespace {

struct BndBuffers {
  vpb_particle_injector_t * send[6] = {}, * recv[6] = {};
  size_t send_cap[6] = {}, recv_cap[6] = {};
  int * d_int = nullptr;     // [0..6) face_count  [8..14) recv counts  [16] n_unknown  [32..32+4*MAX_SP] inject counters
  // particles handed to the host program's boundary handlers, and what the handlers inject
  float4 * custom = nullptr; size_t custom_cap = 0;
  vpb_particle_injector_t * ci = nullptr; size_t ci_cap = 0;
  // sparse STRICT transfers: page-locked host staging and its device twin (rows, then indices)
  char * h_rows = nullptr, * d_rows = nullptr; size_t rows_cap = 0;
  // chained scan of the multi-block kernels: ticket, flags, per-chunk prefixes
  int * d_chain = nullptr; size_t chain_cap = 0;
} B;

// Zeroed scratch for a chained scan over nchunk chunks (on the compute stream).
ChainScan chain_scan( int nchunk ) {
  Runtime & r = rt();
  if( (size_t)nchunk>B.chain_cap ) {
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    if( B.d_chain ) VPB_CUDA( cudaFree( B.d_chain ) );
    B.chain_cap = (size_t)nchunk*2 + 64;
    VPB_CUDA( cudaMalloc( &B.d_chain, sizeof(int)*( 32 + 17*B.chain_cap ) ) );
  }
  VPB_CUDA( cudaMemsetAsync( B.d_chain, 0, sizeof(int)*( 32 + (size_t)nchunk ), r.stream ) );     // ticket + flags
  ChainScan cs; cs.ticket = B.d_chain; cs.flag = B.d_chain + 32; cs.prefix = B.d_chain + 32 + B.chain_cap;
  return cs;
}

void grow( vpb_particle_injector_t ** buf, size_t * cap, size_t want ) {
  if( want<=*cap ) return;
  VPB_CUDA( cudaStreamSynchronize( rt().stream ) );
  if( *buf ) VPB_CUDA( cudaFree( *buf ) );
  *cap = want + want/4 + 256;
  VPB_CUDA( cudaMalloc( buf, sizeof(vpb_particle_injector_t)*( *cap ) ) );
}

// The reference's storage policy before an injection (boundary_p.cc:318-406),
// for species whose arrays live in host memory: grow by 31.25 % when the worst
// case does not fit, shrink when less than half is in use, grow the mover
// array likewise.  The device copies follow the host capacities (mirror()).
constexpr int MIN_NP = 128;
void resize_like_reference( SpeciesMirror * m, int max_inj ) {
  vpb_species_t * sp = m->host;
  if( !sp->p ) return;
  // what the host arrays hold that is worth carrying over: the elements in use, unless the device copy is the
  // current one (RESIDENT), in which case the host array is a stale mirror
  // (STRICT keeps everything: its sparse path patches single rows of the host array afterwards)
  const bool stale = rt().coherence==VPB_COHERENCE_RESIDENT && m->dev_valid && !m->host_valid;
  const size_t keep_p = stale ? 0 : (size_t)sp->max_np, keep_pm = (size_t)sp->max_nm;
  int n = m->np + max_inj;
  if( n>sp->max_np ) {
    n += 0.3125*n;
    VPB_WARNING( "Resizing local %s particle storage from %i to %i", sp->name, sp->max_np, n );
    host_resize_particles( sp, (size_t)n, keep_p );
  } else if( sp->max_np>MIN_NP && n<( sp->max_np>>1 ) ) {
    n += 0.125*n;
    if( n<MIN_NP ) n = MIN_NP;
    VPB_WARNING( "Resizing (shrinking) local %s particle storage from %i to %i", sp->name, sp->max_np, n );
    host_resize_particles( sp, (size_t)n, keep_p );
  }
  int nm = m->nm + max_inj;
  if( nm>sp->max_nm && sp->pm ) {
    nm += 0.3125*nm;
    VPB_WARNING( "This happened.  Resizing local %s mover storage from %i to %i based on not enough movers", sp->name, sp->max_nm, nm );
    host_resize_movers( sp, (size_t)nm, keep_pm );
  }
  mirror( sp );                       // device capacities follow max_np / max_nm
}

// Grow a device-only species' particle storage (the reference reallocates
// sp->p the same way, boundary_p.cc:346-361).
void grow_particles( SpeciesMirror * m, size_t want ) {
  if( want<=m->cap_p ) return;
  vpb_species_t * sp = m->host;
  if( sp->p ) VPB_ERROR( "species \"%s\": %zu particles exceed the device capacity %zu after the host arrays were resized", sp->name, want, m->cap_p );
  Runtime & r = rt();
  size_t n = want + (size_t)( 0.3125*(double)want );
  if( n>(size_t)INT32_MAX ) n = (size_t)INT32_MAX;
  if( n<want ) VPB_ERROR( "species \"%s\": particle count exceeds 32-bit storage", sp->name );
  VPB_WARNING( "Resizing local %s particle storage from %i to %zu", sp->name, sp->max_np, n );
  vpb_particle_t * np_;
  VPB_CUDA( cudaMalloc( &np_, sizeof(vpb_particle_t)*n ) );
  if( m->np>0 ) VPB_CUDA( cudaMemcpyAsync( np_, m->d_p, sizeof(vpb_particle_t)*(size_t)m->np, cudaMemcpyDeviceToDevice, r.stream ) );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  VPB_CUDA( cudaFree( m->d_p ) );
  m->d_p = np_; m->cap_p = n; sp->max_np = (int)n;
}

} // namespace

namespace vpb {
void release_boundary_buffers() {
  for( int f=0; f<6; f++ ) { cudaFree( B.send[f] ); cudaFree( B.recv[f] ); B.send[f] = B.recv[f] = nullptr; B.send_cap[f] = B.recv_cap[f] = 0; }
  cudaFree( B.d_int ); B.d_int = nullptr;
  cudaFree( B.custom ); cudaFree( B.ci ); B.custom = nullptr; B.ci = nullptr; B.custom_cap = B.ci_cap = 0;
  cudaFree( B.d_rows ); if( B.h_rows ) cudaFreeHost( B.h_rows ); B.d_rows = B.h_rows = nullptr; B.rows_cap = 0;
  cudaFree( B.d_chain ); B.d_chain = nullptr; B.chain_cap = 0;
}
}

// ---- STRICT coherence with host arrays: sparse transfers ---------------------
// boundary_p reads the particles its movers name and the tail of the array
// that back-fills their holes, and writes those holes plus what it appends.
// With the particle arrays living in host memory (STRICT) only those rows
// travel; the rest of the device copy is left stale and marked invalid.
struct SparseSpecies { SpeciesMirror * m; int np0, nm0; size_t off;
                       bool was_clean;    // TRACKED coherence: the whole device copy was current on entry (the host did not write the array)
                     };

static void sparse_reserve( size_t rows ) {
  const size_t bytes = rows*( sizeof(vpb_particle_t) + sizeof(int) ) + 256;
  if( bytes<=B.rows_cap ) return;
  VPB_CUDA( cudaStreamSynchronize( rt().stream ) );
  if( B.h_rows ) cudaFreeHost( B.h_rows ); cudaFree( B.d_rows );
  B.rows_cap = bytes + bytes/2;
  VPB_CUDA( cudaMallocHost( &B.h_rows, B.rows_cap ) );
  VPB_CUDA( cudaMalloc( &B.d_rows, B.rows_cap ) );
}

// rows of all species first, then their index lists
static inline vpb_particle_t * rows_at( char * base, size_t off ) { return reinterpret_cast<vpb_particle_t *>( base ) + off; }
static inline int * idx_at( char * base, size_t total, size_t off ) { return reinterpret_cast<int *>( base + total*sizeof(vpb_particle_t) ) + off; }

static void sparse_upload( std::vector<SparseSpecies> & ss ) {
  Runtime & r = rt();
  size_t total = 0;
  for( SparseSpecies & s : ss ) {
    SpeciesMirror * m = s.m; const vpb_species_t * sp = m->host;
    s.was_clean = m->dev_valid && m->np==sp->np && track_clean( sp->p, sizeof(vpb_particle_t)*(size_t)sp->np );
    m->np = sp->np; m->nm = sp->nm; s.np0 = sp->np; s.nm0 = sp->nm; s.off = total; total += (size_t)sp->nm;
    if( m->nm>(int)m->cap_pm ) VPB_ERROR( "species \"%s\": nm=%d exceeds max_nm=%zu", sp->name, m->nm, m->cap_pm );
    if( m->nm>0 && !sp->pm ) VPB_ERROR( "species \"%s\" has movers but no mover array", sp->name );
  }
  sparse_reserve( total );
  for( SparseSpecies & s : ss ) {
    SpeciesMirror * m = s.m; const vpb_species_t * sp = m->host;
    vpb_particle_t * hr = rows_at( B.h_rows, s.off ); int * hi = idx_at( B.h_rows, total, s.off );
    bool sorted = true;
    for( int k=0; k<s.nm0; k++ ) {
      const int i = sp->pm[k].i;
      if( i<0 || i>=s.np0 ) VPB_ERROR( "species \"%s\": mover %d names particle %d of %d", sp->name, k, i, s.np0 );
      if( k && i<=sp->pm[k-1].i ) sorted = false;
      hi[k] = i; hr[k] = sp->p[i];
    }
    m->movers_sorted = sorted;
  }
  if( total ) {
    h2d( B.d_rows, B.h_rows, total*sizeof(vpb_particle_t), r.stream );
    h2d( idx_at( B.d_rows, total, 0 ), idx_at( B.h_rows, total, 0 ), total*sizeof(int), r.stream );
  }
  for( SparseSpecies & s : ss ) {
    SpeciesMirror * m = s.m; const vpb_species_t * sp = m->host;
    if( s.nm0>0 ) {
      h2d( m->d_pm, sp->pm, sizeof(vpb_particle_mover_t)*(size_t)s.nm0, r.stream );
      k_rows_scatter<<< div_up( s.nm0, 256 ), 256, 0, r.stream >>>( m->d_p, idx_at( B.d_rows, total, s.off ), rows_at( B.d_rows, s.off ), s.nm0 );
      r.launches++;
      const int t0 = s.np0-s.nm0>0 ? s.np0-s.nm0 : 0;         // the tail that back-fills the holes
      pin_host( sp->p, sizeof(vpb_particle_t)*(size_t)sp->max_np );
      h2d( m->d_p + t0, sp->p + t0, sizeof(vpb_particle_t)*(size_t)( s.np0-t0 ), r.stream );
    }
    m->dev_valid = true; m->partition_dev_valid = false;
  }
  VPB_CHECK_LAUNCH();
}

static void sparse_download( std::vector<SparseSpecies> & ss ) {
  Runtime & r = rt();
  size_t total = 0;
  for( SparseSpecies & s : ss ) total += (size_t)s.nm0;
  for( SparseSpecies & s : ss ) {
    SpeciesMirror * m = s.m; vpb_species_t * sp = m->host;
    if( s.nm0>0 ) {           // the holes (the index list is still in the device staging buffer)
      k_rows_gather<<< div_up( s.nm0, 256 ), 256, 0, r.stream >>>( rows_at( B.d_rows, s.off ), idx_at( B.d_rows, total, s.off ), m->d_p, s.nm0 );
      r.launches++;
    }
    const int kept = s.np0 - s.nm0;                            // particles appended by the injection
    if( m->np>kept ) d2h( sp->p + kept, m->d_p + kept, sizeof(vpb_particle_t)*(size_t)( m->np-kept ), r.stream );
  }
  VPB_CHECK_LAUNCH();
  if( total ) d2h( B.h_rows, B.d_rows, total*sizeof(vpb_particle_t), r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  for( SparseSpecies & s : ss ) {
    SpeciesMirror * m = s.m; vpb_species_t * sp = m->host;
    const vpb_particle_t * hr = rows_at( B.h_rows, s.off ); const int * hi = idx_at( B.h_rows, total, s.off );
    const int kept = s.np0 - s.nm0;
    for( int k=0; k<s.nm0; k++ ) if( hi[k]<kept ) sp->p[hi[k]] = hr[k];
    sp->np = m->np;
    // only the rows above are current on the device: the host copy is the valid one ...
    m->host_valid = true; m->dev_valid = false; m->partition_dev_valid = false;
    // ... unless the device copy was known to be complete on entry (TRACKED): then both are, and the array
    // (which the patch loop above wrote with CPU stores) is write-protected again
    if( s.was_clean ) { m->dev_valid = true; m->keys_valid = false; track_arm( sp->p, sizeof(vpb_particle_t)*(size_t)sp->max_np ); }
  }
}

// Count phase of the migration (boundary_p.cc:285-300): per axis, our two
// send counts (B.d_int[face]) against the neighbours' (B.d_int[8+face]).
static void exchange_counts( const GridMirror * gm, int * n_recv ) {
  Runtime & r = rt();
  for( int X=0; X<3; X++ ) {
    const int lo = X, hi = X+3;
    const int rank_lo = gm->pkind[lo]==PK_REMOTE ? gm->fbc[lo] : -1, rank_hi = gm->pkind[hi]==PK_REMOTE ? gm->fbc[hi] : -1;
    if( rank_lo<0 && rank_hi<0 ) continue;
    comm_sendrecv( rank_lo, rank_hi, B.d_int+lo, 4, B.d_int+hi, 4, B.d_int+8+lo, 4, B.d_int+8+hi, 4 );
  }
  d2h( n_recv, B.d_int+8, sizeof(int)*6, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
}

extern "C" void boundary_p( void * pbc_list, vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa ) {
  VPB_TRACE_CALL( "boundary_p" );
  if( !sp_list ) return;                                        // boundary_p.cc:77
  if( !fa || !aa || sp_list->g!=aa->g || fa->g!=aa->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  bool counts_done = false; int pre_recv[6] = { 0, 0, 0, 0, 0, 0 };
  {
    // Nothing left the domain and no neighbour rank to talk to: return before
    // touching the mirrors (in STRICT coherence that would re-upload every particle).
    GridMirror * g0 = particle_grid( mirror( sp_list->g ) );
    bool remote = false; long long pending = 0;
    for( int f=0; f<6; f++ ) if( g0->pkind[f]==PK_REMOTE ) remote = true;
    for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) pending += sp->nm;    // always current on the host (movers_to_host)
    if( !remote && pending==0 ) return;
    // Nothing of ours is pending but a neighbour may send: run the count phase
    // of the exchange on its own (all our counts are zero) and stop there when
    // nothing arrives either.  This is what rounds 2 and 3 of a step's
    // num_comm_round look like, and it keeps them from touching any mirror.
    if( pending==0 ) {
      if( !B.d_int ) VPB_CUDA( cudaMalloc( &B.d_int, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ) ) );
      VPB_CUDA( cudaMemsetAsync( B.d_int, 0, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ), r.stream ) );
      exchange_counts( g0, pre_recv );
      counts_done = true;
      int tot = 0; for( int f=0; f<6; f++ ) tot += pre_recv[f];
      if( tot==0 ) return;
    }
  }
  std::vector<SpeciesMirror *> sps;
  bool sparse = r.coherence==VPB_COHERENCE_STRICT;
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    SpeciesMirror * m = mirror( sp );
    if( !sp->p || ( m->dev_valid && !m->host_valid ) ) sparse = false;   // no host array, or the device copy is the current one
  }
  std::vector<SparseSpecies> ss;
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    SpeciesMirror * m = mirror( sp );
    if( sparse ) { SparseSpecies s = { m, 0, 0, 0, false }; ss.push_back( s ); } else to_device( m );
    sps.push_back( m );
    if( sp->id<0 || sp->id>=BND_MAX_SP ) VPB_ERROR( "species id %d is outside the device boundary handler's range", sp->id );
  }
  if( sparse ) sparse_upload( ss );
  // species whose voxel keys (written by advance_p for the lazy sort_p) are worth keeping current
  // (and whose records still have the order of the last sort_p: filling holes and appending keeps the positions'
  // home segments, which is all the segment sort needs)
  std::vector<char> patch( sps.size(), 0 ), home( sps.size(), 0 );
  if( !sparse ) for( size_t k=0; k<sps.size(); k++ ) {
    const SpeciesMirror * m = sps[k];
    patch[k] = ( m->keys_valid || m->keys_patchable ) && m->d_keys && m->cap_keys>=m->cap_p;
    home[k] = m->home_valid;
  }
  GridMirror * gm = particle_grid( sps[0]->grid );
  const GridDev & g = gm->dev;
  bool any_remote = false;
  for( int f=0; f<6; f++ ) if( gm->pkind[f]==PK_REMOTE ) any_remote = true;
  long long nm_tot = 0;
  for( SpeciesMirror * m : sps ) nm_tot += m->nm;
  if( !any_remote && nm_tot==0 ) { if( sparse ) sparse_download( ss ); return; }   // nothing left the domain and nobody to talk to
  // User handlers (boundary_p.cc:81-89): indexed by -id-3.  They are host
  // functions of the host program, called below in the reference's order.
  constexpr int MAX_PBC = 32;
  vpb_particle_bc_func_t pbc_interact[MAX_PBC] = {}; void * pbc_params[MAX_PBC] = {};
  bool any_custom = false;
  unsigned handler_mask = 0;
  for( const vpb_particle_bc_t * pbc=(const vpb_particle_bc_t *)pbc_list; pbc; pbc=pbc->next ) {
    const long long k = -pbc->id-3;
    if( k<0 || k>=MAX_PBC ) VPB_ERROR( "Update this to support more particle boundary conditions" );
    pbc_interact[k] = pbc->interact; pbc_params[k] = pbc->params;
    if( pbc->interact ) handler_mask |= 1u<<k;
  }
  if( gm->per_voxel && gm->any_custom && handler_mask ) any_custom = true;
  for( int f=0; f<6; f++ ) if( gm->pkind[f]==PK_CUSTOM && gm->pcode[f]<=-3 ) {
    const long long k = -gm->pcode[f]-3;
    if( k<MAX_PBC && pbc_interact[k] ) any_custom = true;
  }
  FieldMirror * mf = mirror( fa );
  AccMirror   * ma = mirror( aa );
  bool any_absorb = false;
  for( int f=0; f<6; f++ ) if( gm->pkind[f]==PK_ABSORB ) any_absorb = true;
  if( gm->per_voxel && gm->any_absorb ) any_absorb = true;
  if( any_absorb || !sparse ) to_device( mf );      // rhob is only touched by absorbed particles
  if( !sparse ) to_device( ma );                    // (sparse: uploaded when something arrives)
  if( !B.d_int ) VPB_CUDA( cudaMalloc( &B.d_int, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ) ) );
  VPB_CUDA( cudaMemsetAsync( B.d_int, 0, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ), r.stream ) );

  // ---- load the send buffers, absorb, back-fill (boundary_p.cc:130-283) ----
  BndFaces bf;
  for( int f=0; f<6; f++ ) {
    bf.kind[f] = gm->pkind[f]; bf.nb_n[f] = gm->nb_n[f]; bf.code[f] = (int)gm->pcode[f];
    if( gm->pkind[f]==PK_REMOTE ) grow( &B.send[f], &B.send_cap[f], (size_t)nm_tot );
    bf.send[f] = B.send[f];
  }
  bf.face_count = B.d_int; bf.handler_mask = handler_mask;
  bf.custom = nullptr;
  if( any_custom && nm_tot>0 ) {
    if( (size_t)nm_tot>B.custom_cap ) {
      VPB_CUDA( cudaStreamSynchronize( r.stream ) );
      if( B.custom ) VPB_CUDA( cudaFree( B.custom ) );
      B.custom_cap = (size_t)nm_tot + (size_t)nm_tot/4 + 256;
      VPB_CUDA( cudaMalloc( &B.custom, sizeof(float4)*4*B.custom_cap ) );
    }
    bf.custom = B.custom;
  }
  bool touched_rhob = false;
  for( size_t ks=0; ks<sps.size(); ks++ ) {
    SpeciesMirror * m = sps[ks];
    if( m->nm<=0 ) continue;
    if( !m->movers_sorted ) { sort_movers_by_index( m ); m->movers_sorted = true; }
    const vpb_species_t * sp = m->host;
    const float q_r8V = sp->q*sp->g->r8V;
    if( r.deposit_mode==VPB_DEPOSIT_PARITY || m->nm<=BND_T )
      k_bnd_classify<<< 1, BND_T, 0, r.stream >>>( m->d_p, m->d_pm, m->nm, sp->id, q_r8V, mf->comp( F_RHOB ), g, bf, B.d_int+16,
                                                    r.deposit_mode==VPB_DEPOSIT_PARITY );
    else {
      const int nchunk = div_up( m->nm, BND_T );
      k_bnd_classify_mb<<< nchunk, BND_T, 0, r.stream >>>( m->d_p, m->d_pm, m->nm, sp->id, q_r8V, mf->comp( F_RHOB ), g, bf, B.d_int+16,
                                                            chain_scan( nchunk ), nchunk );
    }
    k_bnd_backfill<<< div_up( m->nm, 256 ), 256, 0, r.stream >>>( m->d_p, m->d_pm, m->nm, m->np );
    VPB_CHECK_LAUNCH(); r.launches += 2;
    if( patch[ks] ) { k_keys_patch<<< div_up( m->nm, 256 ), 256, 0, r.stream >>>( m->d_keys, m->d_p, m->d_pm, m->nm, 0, m->np-m->nm ); r.launches++; }
    if( any_absorb ) touched_rhob = true;
    m->np -= m->nm; m->nm = 0; m->partition_dev_valid = false;
    mark_dev_written( m );
  }
  int h_int[32];
  d2h( h_int, B.d_int, sizeof(int)*32, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( h_int[16] ) VPB_WARNING( "Unknown boundary interaction ... dropping %d particles", h_int[16] );   // boundary_p.cc:259-260
  int n_send[6], n_recv[6] = { 0, 0, 0, 0, 0, 0 };
  for( int f=0; f<6; f++ ) n_send[f] = h_int[f];

  // ---- user handlers (boundary_p.cc:236-253) ----
  // The particles that hit a handler's face come back in processing order
  // (species in list order, movers last-first); each is given to the host
  // program's interact() with its own copy of the particle and mover, and the
  // injectors it returns form the local list that is injected first.
  int n_ci = 0;
  const int n_custom = h_int[6];
  if( n_custom>0 ) {
    std::vector<float4> ent( 4*(size_t)n_custom );
    d2h( ent.data(), B.custom, sizeof(float4)*4*(size_t)n_custom, r.stream );
    // handlers may read or update the host fields (absorb_tally -> accumulate_rhob):
    // the charge absorbed on the device goes down first
    if( touched_rhob && r.coherence==VPB_COHERENCE_STRICT ) { mark_dev_written( mf ); to_host( mf ); touched_rhob = false; }
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    std::vector<vpb_particle_injector_t> ci( (size_t)n_custom );
    vpb_species_t * sp_by_id[BND_MAX_SP] = {};
    for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) sp_by_id[sp->id] = sp;
    for( int k=0; k<n_custom; k++ ) {
      vpb_particle_t prt; vpb_particle_mover_t mv;
      memcpy( &prt, &ent[4*(size_t)k], sizeof(prt) );
      memcpy( &mv, &ent[4*(size_t)k+2], sizeof(mv) );
      int face, sp_id, code;
      memcpy( &face, &ent[4*(size_t)k+3].x, 4 ); memcpy( &sp_id, &ent[4*(size_t)k+3].y, 4 ); memcpy( &code, &ent[4*(size_t)k+3].z, 4 );
      const long long h = -(long long)code-3;
      if( (size_t)n_ci+1>ci.size() ) ci.resize( ci.size()*2 );
      n_ci += pbc_interact[h]( pbc_params[h], sp_by_id[sp_id], &prt, &mv, ci.data()+n_ci, 1, face );
    }
    if( n_ci>0 ) {
      grow( &B.ci, &B.ci_cap, (size_t)n_ci );
      h2d( B.ci, ci.data(), sizeof(vpb_particle_injector_t)*(size_t)n_ci, r.stream );
      VPB_CUDA( cudaStreamSynchronize( r.stream ) );      // ci is a local vector
    }
    if( r.coherence==VPB_COHERENCE_STRICT ) { mf->dev_valid = false; mf->host_valid = true; }   // a handler may have written host fields
  }

  // ---- exchange counts, then injectors (boundary_p.cc:285-316) ----
  if( any_remote ) {
    if( counts_done ) for( int f=0; f<6; f++ ) n_recv[f] = pre_recv[f];
    else exchange_counts( gm, n_recv );
    for( int f=0; f<6; f++ ) if( gm->pkind[f]==PK_REMOTE ) grow( &B.recv[f], &B.recv_cap[f], (size_t)n_recv[f] ); else n_recv[f] = 0;
    const size_t S = sizeof(vpb_particle_injector_t);
    for( int X=0; X<3; X++ ) {
      const int lo = X, hi = X+3;
      int rank_lo = gm->pkind[lo]==PK_REMOTE ? gm->fbc[lo] : -1, rank_hi = gm->pkind[hi]==PK_REMOTE ? gm->fbc[hi] : -1;
      // Whether a pair of neighbours exchanges a payload is decided from the two
      // counts BOTH of them hold after the count phase (ours to them, theirs to
      // us), never from traffic on our other face: with three or more ranks along
      // an axis the rank on the quiet side has no way to know about that.
      if( n_send[lo]+n_recv[lo]==0 ) rank_lo = -1;
      if( n_send[hi]+n_recv[hi]==0 ) rank_hi = -1;
      if( rank_lo<0 && rank_hi<0 ) continue;
      comm_sendrecv( rank_lo, rank_hi, B.send[lo], S*(size_t)n_send[lo], B.send[hi], S*(size_t)n_send[hi],
                     B.recv[lo], S*(size_t)n_recv[lo], B.recv[hi], S*(size_t)n_recv[hi] );
    }
  }

  // ---- inject (boundary_p.cc:318-494) ----
  int max_inj = n_ci;
  for( int f=0; f<6; f++ ) max_inj += n_recv[f];
  for( SpeciesMirror * m : sps ) resize_like_reference( m, max_inj );
  if( max_inj>0 ) {
    if( sparse ) to_device( ma );
    BndInject bi = {};
    bi.start[0] = 0;
    bi.recv[0] = B.ci; bi.start[1] = n_ci;                                  // custom local injection first (:426-432)
    for( int f=0; f<6; f++ ) { bi.recv[f+1] = B.recv[f]; bi.start[f+2] = bi.start[f+1] + n_recv[f]; }
    for( SpeciesMirror * m : sps ) {
      grow_particles( m, (size_t)m->np + (size_t)max_inj );
      if( (size_t)m->nm + (size_t)max_inj>m->cap_pm && !m->host->pm ) VPB_ERROR( "species \"%s\": raise max_nm", m->host->name );
      const int id = m->host->id;
      bi.p[id] = m->d_p; bi.pm[id] = m->d_pm; bi.np[id] = m->np; bi.nm[id] = m->nm; bi.max_nm[id] = (int)m->cap_pm;
      bi.q[id] = m->host->q;
    }
    bi.counters = B.d_int + 32;
    if( r.deposit_mode==VPB_DEPOSIT_PARITY ) k_bnd_inject_serial<<< 1, 32, 0, r.stream >>>( bi, ma->d, g );
    else if( max_inj<=BND_T ) k_bnd_inject<<< 1, BND_T, 0, r.stream >>>( bi, ma->d, g );
    else {
      const int nchunk = div_up( max_inj, BND_T );
      k_bnd_inject_mb<<< nchunk, BND_T, 0, r.stream >>>( bi, ma->d, g, chain_scan( nchunk ), nchunk );
    }
    VPB_CHECK_LAUNCH(); r.launches++;
    int h_cnt[4*BND_MAX_SP+1];
    d2h( h_cnt, B.d_int+32, sizeof(h_cnt), r.stream );
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    if( h_cnt[3*BND_MAX_SP] ) VPB_WARNING( "Dropped %d received particles of unknown species", h_cnt[3*BND_MAX_SP] );
    for( size_t ks=0; ks<sps.size(); ks++ ) {
      SpeciesMirror * m = sps[ks];
      const int id = m->host->id;
      if( patch[ks] && h_cnt[id]>0 ) {
        if( m->cap_keys<m->cap_p ) patch[ks] = 0;            // the particle array grew past the key array
        else { k_keys_patch<<< div_up( h_cnt[id], 256 ), 256, 0, r.stream >>>( m->d_keys, m->d_p, nullptr, h_cnt[id], m->np, m->np+h_cnt[id] ); r.launches++; }
      }
      m->np += h_cnt[id];
      const int app = h_cnt[BND_MAX_SP+id], drop = h_cnt[2*BND_MAX_SP+id];
      if( drop ) patch[ks] = 0;                              // a particle stuck with 8*voxel+face and no mover
      if( drop ) VPB_WARNING( "%i particles were not completed moved to their final location this timestep for species \"%s\"", drop, m->host->name );
      m->nm += app - drop;
      if( app-drop>0 ) m->movers_sorted = false;
      if( h_cnt[id] ) { m->partition_dev_valid = false; mark_dev_written( m ); }
    }
    mark_dev_written( ma );
  }
  if( touched_rhob ) mark_dev_written( mf );

  VPB_CHECK_LAUNCH();
  for( size_t ks=0; ks<sps.size(); ks++ ) {
    SpeciesMirror * m = sps[ks];
    if( patch[ks] ) { m->keys_valid = m->nm==0; m->keys_patchable = m->nm>0; }
    if( home[ks] ) m->home_valid = true;
    m->host->np = m->np;
    if( m->nm>1 && !m->movers_sorted ) { sort_movers_by_index( m ); m->movers_sorted = true; }
    movers_to_host( m );
    if( r.coherence==VPB_COHERENCE_STRICT && !sparse ) to_host( m );
  }
  if( sparse ) sparse_download( ss );
  if( r.coherence==VPB_COHERENCE_STRICT ) { to_host( ma ); to_host( mf ); }
}

// What vpic_simulation::advance() does with movers that are still unprocessed
// after num_comm_round calls of boundary_p (src/vpic/advance.cc:78-101): the
// particle's voxel is restored (p.i >>= 3), its charge goes to rhob
// (accumulate_rhob), it is removed with the usual back-fill, last mover first.
// A host program built from the reference runs that loop itself on the host
// arrays; this is the same thing for programs whose particles live on the
// device (the stand-alone driver, bench.py).  Returns the particles removed.
extern "C" int vpic_b200_drop_unprocessed_movers( vpb_species_t * sp_list, vpb_field_array_t * fa ) {
  if( !sp_list ) return 0;
  if( !fa || sp_list->g!=fa->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  int dropped = 0;
  FieldMirror * mf = nullptr;
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    if( sp->nm<=0 ) continue;
    SpeciesMirror * m = mirror( sp );
    to_device( m );
    if( m->nm<=0 ) continue;
    if( !mf ) { mf = mirror( fa ); to_device( mf ); }
    GridMirror * gm = particle_grid( m->grid );
    GridDev g = gm->dev; g.neighbor = nullptr;                    // every exit is absorbed, whatever the table says
    if( !B.d_int ) VPB_CUDA( cudaMalloc( &B.d_int, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ) ) );
    VPB_CUDA( cudaMemsetAsync( B.d_int, 0, sizeof(int)*( 32 + 4*BND_MAX_SP + 8 ), r.stream ) );
    BndFaces bf = {};
    for( int f=0; f<6; f++ ) bf.kind[f] = PK_ABSORB;
    bf.face_count = B.d_int;
    if( !m->movers_sorted ) { sort_movers_by_index( m ); m->movers_sorted = true; }
    VPB_WARNING( "Removing %i particles associated with unprocessed %s movers (increase num_comm_round)", m->nm, sp->name );
    const float q_r8V = sp->q*sp->g->r8V;
    // single ordered block: the movers are few by construction, and rhob receives
    // them one after the other like the host loop
    k_bnd_classify<<< 1, BND_T, 0, r.stream >>>( m->d_p, m->d_pm, m->nm, sp->id, q_r8V, mf->comp( F_RHOB ), g, bf, B.d_int+16, true );
    k_bnd_backfill<<< div_up( m->nm, 256 ), 256, 0, r.stream >>>( m->d_p, m->d_pm, m->nm, m->np );
    VPB_CHECK_LAUNCH(); r.launches += 2;
    const bool patch = m->keys_patchable && m->d_keys && m->cap_keys>=m->cap_p, home = m->home_valid;
    if( patch ) { k_keys_patch<<< div_up( m->nm, 256 ), 256, 0, r.stream >>>( m->d_keys, m->d_p, m->d_pm, m->nm, 0, m->np-m->nm ); VPB_CHECK_LAUNCH(); r.launches++; }
    dropped += m->nm;
    m->np -= m->nm; m->nm = 0; m->partition_dev_valid = false;
    mark_dev_written( m );
    m->keys_valid = patch; m->home_valid = home;
    sp->np = m->np;
    movers_to_host( m );
    if( r.coherence==VPB_COHERENCE_STRICT ) to_host( m );
  }
  if( mf ) finish_write( mf );
  return dropped;
}
