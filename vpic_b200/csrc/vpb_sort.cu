// vpb_sort.cu -- sort_p: stable counting sort of a species by voxel.
// Reference: src/species_advance/standard/pipeline/sort_p_pipeline.c:143-371
// (the result of its single-pipeline subsort is what we reproduce exactly:
// particles grouped by voxel in ascending voxel order, ORIGINAL ORDER KEPT
// inside each voxel, partition[v] = first index of voxel v).
//
// Device algorithms (all hand-written, no CUB); which one runs is decided per call in sort_p() below:
//
//   generic, immediate (STRICT coherence or PARITY deposit)
//     1. keys[i] = p[i].i and a voxel histogram (warp-aggregated atomics),
//     2. exclusive scan of the histogram -> partition[],
//     3. scatter the particle INDICES into their voxel's segment (atomic slots;
//        order inside a segment is arbitrary at this point),
//     4. the warp that owns a voxel ranks its segment ascending (this is what
//        makes the sort stable and run-to-run deterministic) and, in the same
//        kernel, copies the 32 B particles to their final places in the spare
//        buffer (2 x 128-bit loads / stores each); buffers are swapped.
//   generic, lazy (RESIDENT coherence + FAST deposit)
//     the same, but step 4 leaves the ordered indices in d_perm and copies nothing:
//     the next advance_p gathers through them (k_advance_p<GATHER>); step 1 reads
//     the 4-byte keys the previous advance_p left when they are there.
//   segment sort (lazy, from a species' second sort on)
//     deterministic ranks from the order the previous sort left: see the comment
//     above k_seg_pass.
//
// Algorithmic traffic: 4 B key + 32 B read + 32 B write per particle (SURVEY
// 8d: 68 B).  The immediate sort moves ~120 B because of the index passes; the
// lazy ones move 20 - 25 B themselves and leave the 64 B copy to the push.
#include "vpb_runtime.h"
#include "vpb_sortutil.cuh"

namespace vpb {

// ---- generic exclusive scan of n ints (3 phases) ---------------------------
constexpr int SCAN_T = 256, SCAN_E = 16, SCAN_TILE = SCAN_T*SCAN_E;

__global__ void __launch_bounds__(SCAN_T)
k_scan_tiles( const int * __restrict__ in, int * __restrict__ out, int * __restrict__ tile_sum, int n ) {
  __shared__ int s_warp[SCAN_T/32];
  const long long base = (long long)blockIdx.x*SCAN_TILE + (long long)threadIdx.x*SCAN_E;
  int v[SCAN_E], sum = 0;
# pragma unroll
  for( int k=0; k<SCAN_E; k++ ) { v[k] = base+k<n ? in[base+k] : 0; sum += v[k]; }
  // inclusive warp scan of per-thread sums
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = sum;
  for( int o=1; o<32; o<<=1 ) { int t = __shfl_up_sync( 0xffffffffu, inc, o ); if( lane>=o ) inc += t; }
  if( lane==31 ) s_warp[w] = inc;
  __syncthreads();
  if( w==0 ) {
    int t = lane<SCAN_T/32 ? s_warp[lane] : 0, i2 = t;
    for( int o=1; o<32; o<<=1 ) { int u = __shfl_up_sync( 0xffffffffu, i2, o ); if( lane>=o ) i2 += u; }
    if( lane<SCAN_T/32 ) s_warp[lane] = i2 - t;             // exclusive warp offsets
    if( lane==SCAN_T/32-1 ) tile_sum[blockIdx.x] = i2;
  }
  __syncthreads();
  int run = s_warp[w] + inc - sum;
# pragma unroll
  for( int k=0; k<SCAN_E; k++ ) { if( base+k<n ) out[base+k] = run; run += v[k]; }
}

__global__ void k_scan_sums( int * __restrict__ tile_sum, int ntile ) {   // single block, sequential chunks
  __shared__ int s_carry; __shared__ int s_warp[32];
  if( threadIdx.x==0 ) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for( int base=0; base<ntile; base+=blockDim.x ) {
    const int i = base + threadIdx.x;
    const int v = i<ntile ? tile_sum[i] : 0;
    int inc = v;
    for( int o=1; o<32; o<<=1 ) { int t = __shfl_up_sync( 0xffffffffu, inc, o ); if( lane>=o ) inc += t; }
    if( lane==31 ) s_warp[w] = inc;
    __syncthreads();
    if( w==0 ) {
      int t = lane<(int)(blockDim.x>>5) ? s_warp[lane] : 0, i2 = t;
      for( int o=1; o<32; o<<=1 ) { int u = __shfl_up_sync( 0xffffffffu, i2, o ); if( lane>=o ) i2 += u; }
      s_warp[lane] = i2 - t;
    }
    __syncthreads();
    const int carry = s_carry;
    if( i<ntile ) tile_sum[i] = carry + s_warp[w] + inc - v;
    __syncthreads();
    if( threadIdx.x==blockDim.x-1 ) s_carry = carry + s_warp[w] + inc;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SCAN_T)
k_scan_add( int * __restrict__ out, int * __restrict__ out2, const int * __restrict__ tile_sum, int n ) {
  const int add = tile_sum[blockIdx.x];
  const long long base = (long long)blockIdx.x*SCAN_TILE;
  for( int k=threadIdx.x; k<SCAN_TILE; k+=SCAN_T ) if( base+k<n ) { int v = out[base+k] + add; out[base+k] = v; if( out2 ) out2[base+k] = v; }
}

// exclusive scan in -> out (and a second copy out2 if not null); tmp >= div_up(n,SCAN_TILE) ints
void exclusive_scan( const int * in, int * out, int * out2, int * tmp, int n ) {
  Runtime & r = rt();
  const int ntile = div_up( n, SCAN_TILE );
  k_scan_tiles<<< ntile, SCAN_T, 0, r.stream >>>( in, out, tmp, n );
  k_scan_sums<<< 1, 1024, 0, r.stream >>>( tmp, ntile );
  k_scan_add<<< ntile, SCAN_T, 0, r.stream >>>( out, out2, tmp, n );
  VPB_CHECK_LAUNCH(); r.launches += 3;
}

// ---- sort kernels ----------------------------------------------------------

// One thread per particle and chunk, SORT_CHUNKS chunks of 256 consecutive particles per block: with one chunk
// per block the 262 144 tiny blocks of a 67 M particle species were bound by the block launch rate (228 us for
// the key histogram against 45 us of traffic), not by their atomics.
constexpr int SORT_CHUNKS = 16;

// The kernels are latency bound (one key load, one atomic round trip per warp and chunk at full occupancy: 40 warps
// stalled on the scoreboard per issue), so every thread works on SORT_ILP chunks at once: the loads, the matches
// and the atomics of the batch are issued back to back before the first result is needed.
constexpr int SORT_ILP = 4;
static_assert( SORT_CHUNKS % SORT_ILP==0, "chunks per block come in batches" );

template<bool FROM_PARTICLES>
__device__ __forceinline__ void hist_block( const vpb_particle_t * __restrict__ p, const int * __restrict__ keys_in, int np,
                                            int * __restrict__ keys_out, int * __restrict__ count ) {
  const int lane = threadIdx.x & 31;
  for( int c0=0; c0<SORT_CHUNKS; c0+=SORT_ILP ) {
    const long long i0 = ( (long long)blockIdx.x*SORT_CHUNKS + c0 )*256 + threadIdx.x;
    if( i0 - threadIdx.x>=np ) return;
    int key[SORT_ILP]; unsigned m[SORT_ILP];
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) { const long long i = i0 + 256*u; key[u] = i<np ? ( FROM_PARTICLES ? p[i].i : keys_in[i] ) : -1; }
    if( FROM_PARTICLES ) {
#     pragma unroll
      for( int u=0; u<SORT_ILP; u++ ) if( key[u]>=0 ) keys_out[i0 + 256*u] = key[u];
    }
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) m[u] = __match_any_sync( 0xffffffffu, key[u] );
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) if( key[u]>=0 && lane==__ffs( m[u] )-1 ) atomicAdd( count + key[u], __popc( m[u] ) );
  }
}

__global__ void __launch_bounds__(256)
k_sort_hist( const vpb_particle_t * __restrict__ p, int np, int * __restrict__ keys, int * __restrict__ count ) {
  hist_block<true>( p, nullptr, np, keys, count );
}

__global__ void __launch_bounds__(256)
k_sort_scatter_idx( const int * __restrict__ keys, int np, int * __restrict__ next, int * __restrict__ perm ) {
  const int lane = threadIdx.x & 31;
  for( int c0=0; c0<SORT_CHUNKS; c0+=SORT_ILP ) {
    const long long i0 = ( (long long)blockIdx.x*SORT_CHUNKS + c0 )*256 + threadIdx.x;
    if( i0 - threadIdx.x>=np ) return;
    int key[SORT_ILP], base[SORT_ILP]; unsigned m[SORT_ILP];
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) { const long long i = i0 + 256*u; key[u] = i<np ? keys[i] : -1; }
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) m[u] = __match_any_sync( 0xffffffffu, key[u] );
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) { base[u] = 0; if( key[u]>=0 && lane==__ffs( m[u] )-1 ) base[u] = atomicAdd( next + key[u], __popc( m[u] ) ); }
#   pragma unroll
    for( int u=0; u<SORT_ILP; u++ ) {
      base[u] = __shfl_sync( 0xffffffffu, base[u], __ffs( m[u] )-1 );
      if( key[u]>=0 ) perm[ base[u] + __popc( m[u] & ( ( 1u<<lane )-1u ) ) ] = (int)( i0 + 256*u );
    }
  }
}

// Ascending sort of every voxel's segment of perm[].  One warp per voxel.
// Segments <= 32: shuffle bitonic.  <= FIX_SMEM: bitonic in shared memory.
// Larger: bitonic in global memory by the same warp (rare: > 2048 particles in
// one voxel).

__device__ __forceinline__ void copy_particle( vpb_particle_t * __restrict__ out, int dst, const vpb_particle_t * __restrict__ in, int src ) {
  const float4 * s = reinterpret_cast<const float4 *>( in + src );
  const float4 a = __ldg( s ), b = __ldg( s+1 );
  float4 * d = reinterpret_cast<float4 *>( out + dst );
  d[0] = a; d[1] = b;
}

// Fix-up and gather fused: the warp that owns voxel v puts the particle
// indices of its segment in ascending order (= original array order: that is
// what makes the sort stable) and copies the particles to their final places.
// Segments of <= 32*FIX_E indices sort in registers (a bitonic network of 32,
// 64 or 128 elements, chosen per segment), larger ones fall back to the
// bitonic sorts in shared or global memory.
constexpr int FIX_E = 4, FIX_REP = 8;

// Bitonic sort of 32*E ints held E per lane (element index 32k+lane): strides
// of 32 and up exchange registers of one lane, shorter strides use shuffles.
template<int E> __device__ __forceinline__ void warp_bitonic_regs( int (&x)[E], int lane ) {
# pragma unroll
  for( int sz=2; sz<=32*E; sz<<=1 ) {
#   pragma unroll
    for( int j=sz>>1; j>0; j>>=1 ) {
      if( j>=32 ) {
        const int dk = j>>5;
#       pragma unroll
        for( int k=0; k<E; k++ ) if( ( k & dk )==0 ) {
          const bool up = ( ( 32*k ) & sz )==0;
          const int lo = min( x[k], x[k|dk] ), hi = max( x[k], x[k|dk] );
          x[k] = up ? lo : hi; x[k|dk] = up ? hi : lo;
        }
      } else {
        const bool lower = ( lane & j )==0;
#       pragma unroll
        for( int k=0; k<E; k++ ) {
          const bool up = ( ( 32*k+lane ) & sz )==0;
          const int y = __shfl_xor_sync( 0xffffffffu, x[k], j );
          x[k] = ( lower==up ) ? min( x[k], y ) : max( x[k], y );
        }
      }
    }
  }
}

// GATHER: copy the particles to their final places; otherwise (lazy sort_p) put the segment of perm[] in order
// IN PLACE -- the next advance_p gathers through it -- which for a segment that arrived in order is no write at all.
template<int E, bool GATHER> __device__ __forceinline__ void
fix_gather_regs( int * perm, int b, int n, int lane, const vpb_particle_t * __restrict__ in, vpb_particle_t * __restrict__ out ) {
  int x[E];
# pragma unroll
  for( int k=0; k<E; k++ ) x[k] = 32*k+lane<n ? perm[b+32*k+lane] : 0x7fffffff;
  bool ok = true;
# pragma unroll
  for( int k=0; k<E; k++ ) {
    const int nxt = __shfl_down_sync( 0xffffffffu, x[k], 1 );
    const int wrap = __shfl_sync( 0xffffffffu, k+1<E ? x[k+1<E ? k+1 : k] : 0x7fffffff, 0 );
    ok = ok && x[k]<=( lane==31 ? wrap : nxt );
  }
  const bool sorted = __all_sync( 0xffffffffu, ok );
  if( !sorted ) warp_bitonic_regs<E>( x, lane );
# pragma unroll
  for( int k=0; k<E; k++ ) if( 32*k+lane<n ) { if( GATHER ) copy_particle( out, b+32*k+lane, in, x[k] ); else if( !sorted ) perm[b+32*k+lane] = x[k]; }
}

// A warp owns FIX_REP consecutive voxels.  The kernel was latency bound (partition load -> segment load -> store,
// one voxel after the other, 34 warps stalled per issue), so the common case -- every one of the warp's segments
// holds at most 64 indices -- loads all of them before it looks at the first.  The atomic slots of the scatter
// went to whole warps of a nearly sorted array, so most segments arrive ascending already: one neighbour compare
// per element finds out and skips the sorting network.
template<bool GATHER>
__global__ void __launch_bounds__(32*FIX_WARPS)
k_sort_fix_gather( const int * __restrict__ partition, int nv, int np, int * perm, int * __restrict__ big_scratch,
                   const vpb_particle_t * __restrict__ in, vpb_particle_t * __restrict__ out ) {
  __shared__ int s[FIX_WARPS][FIX_SMEM];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int v0 = ( blockIdx.x*FIX_WARPS + w )*FIX_REP;
  if( v0>=nv ) return;
  const int vl = v0 + lane;
  const int pl = ( lane<=FIX_REP && vl<nv ) ? partition[vl] : np;     // partition[v0 .. v0+FIX_REP], np past the end
  int b[FIX_REP], n[FIX_REP], nmax = 0;
# pragma unroll
  for( int r=0; r<FIX_REP; r++ ) {
    b[r] = __shfl_sync( 0xffffffffu, pl, r ); n[r] = __shfl_sync( 0xffffffffu, pl, r+1 ) - b[r];
    nmax = max( nmax, n[r] );
  }
  if( nmax<=64 ) {
    int x[FIX_REP][2];
#   pragma unroll
    for( int r=0; r<FIX_REP; r++ ) {
      x[r][0] = lane<n[r]    ? perm[b[r]+lane]    : 0x7fffffff;
      x[r][1] = 32+lane<n[r] ? perm[b[r]+32+lane] : 0x7fffffff;
    }
#   pragma unroll
    for( int r=0; r<FIX_REP; r++ ) {
      if( n[r]<=0 ) continue;
      const int nxt0 = __shfl_down_sync( 0xffffffffu, x[r][0], 1 ), nxt1 = __shfl_down_sync( 0xffffffffu, x[r][1], 1 );
      const int wrap = __shfl_sync( 0xffffffffu, x[r][1], 0 );
      const bool ok = x[r][0]<=( lane==31 ? wrap : nxt0 ) && ( lane==31 || x[r][1]<=nxt1 );
      const bool sorted = __all_sync( 0xffffffffu, ok );
      if( !sorted ) {
        if( n[r]<=32 ) { int y[1] = { x[r][0] }; warp_bitonic_regs<1>( y, lane ); x[r][0] = y[0]; }
        else warp_bitonic_regs<2>( x[r], lane );
      }
      if( GATHER ) {
        if( lane<n[r] )    copy_particle( out, b[r]+lane, in, x[r][0] );
        if( 32+lane<n[r] ) copy_particle( out, b[r]+32+lane, in, x[r][1] );
      } else if( !sorted ) {
        if( lane<n[r] )    perm[b[r]+lane] = x[r][0];
        if( 32+lane<n[r] ) perm[b[r]+32+lane] = x[r][1];
      }
    }
    return;
  }
  // some segment is longer: one voxel after the other; <= 32*FIX_E indices sort in registers, longer ones in
  // shared memory, > FIX_SMEM in global memory (rare: a crowded voxel)
  for( int r=0; r<FIX_REP; r++ ) {
    const int bb = __shfl_sync( 0xffffffffu, pl, r ), nn = __shfl_sync( 0xffffffffu, pl, r+1 ) - bb;
    if( nn<=0 ) continue;
    if( nn<=32 )       { fix_gather_regs<1,GATHER>( perm, bb, nn, lane, in, out ); continue; }
    if( nn<=64 )       { fix_gather_regs<2,GATHER>( perm, bb, nn, lane, in, out ); continue; }
    if( nn<=32*FIX_E ) { fix_gather_regs<FIX_E,GATHER>( perm, bb, nn, lane, in, out ); continue; }
    int n2 = 64; while( n2<nn ) n2 <<= 1;
    int * a;
    if( n2<=FIX_SMEM ) a = s[w];
    else a = big_scratch + 2*(size_t)bb;     // window of n2 < 2n ints starting at 2b: windows of different voxels are disjoint
    for( int t=lane; t<n2; t+=32 ) a[t] = t<nn ? perm[bb+t] : 0x7fffffff;
    __threadfence_block(); __syncwarp();
    bitonic_mem( a, n2, lane );
    for( int t=lane; t<nn; t+=32 ) { if( GATHER ) copy_particle( out, bb+t, in, a[t] ); else perm[bb+t] = a[t]; }
    __syncwarp();
  }
}

// out[d] = in[ perm[d] ]: carries out a lazy sort that no advance_p consumed
__global__ void __launch_bounds__(256)
k_sort_apply_perm( const int * __restrict__ perm, int np, const vpb_particle_t * __restrict__ in, vpb_particle_t * __restrict__ out ) {
  const int d = blockIdx.x*blockDim.x + threadIdx.x;
  if( d<np ) copy_particle( out, d, in, perm[d] );
}

__global__ void __launch_bounds__(256)
k_hist_keys( const int * __restrict__ keys, int n, int * __restrict__ count ) {
  hist_block<false>( nullptr, keys, n, nullptr, count );
}

void bin_indices( const int * keys, int n, int nbin, int * start, int * perm ) {
  Runtime & r = rt();
  const size_t nba = ( (size_t)nbin + 1 + 31 ) & ~(size_t)31;
  const size_t ntile = (size_t)div_up( nbin+1, SCAN_TILE ) + 32;
  int * base = (int *)scratch( sizeof(int)*( 2*nba + ntile + 64 ) );
  int * count = base, * next = count + nba, * tiles = next + nba;
  VPB_CUDA( cudaMemsetAsync( count, 0, sizeof(int)*( (size_t)nbin+1 ), r.stream ) );
  if( n>0 ) { k_hist_keys<<< div_up(n,256*SORT_CHUNKS), 256, 0, r.stream >>>( keys, n, count ); r.launches++; }
  exclusive_scan( count, start, next, tiles, nbin+1 );
  if( n>0 ) { k_sort_scatter_idx<<< div_up(n,256*SORT_CHUNKS), 256, 0, r.stream >>>( keys, n, next, perm ); r.launches++; }
  VPB_CHECK_LAUNCH();
}

// Diagnostic for tests at sizes no CPU oracle reaches: particles out of voxel
// order, or outside the segment partition[] gives their voxel.
__global__ void __launch_bounds__(256)
k_sort_check( const vpb_particle_t * __restrict__ p, int np, const int * __restrict__ partition, int nv, unsigned long long * __restrict__ bad ) {
  const int i = blockIdx.x*blockDim.x + threadIdx.x;
  if( i>=np ) return;
  const int key = p[i].i;
  bool wrong = key<0 || key>=nv;
  if( !wrong ) {
    const int b = partition[key], e = key+1<nv ? partition[key+1] : np;
    wrong = i<b || i>=e || ( i>0 && p[i-1].i>key );
  }
  if( wrong ) atomicAdd( bad, 1ull );
}

// spare particle buffer shared by all species (swapped with the sorted species)
static vpb_particle_t * g_alt = nullptr; static size_t g_alt_cap = 0;
void release_sort_buffers() { if( g_alt ) cudaFree( g_alt ); g_alt = nullptr; g_alt_cap = 0; }

vpb_particle_t * sort_spare( size_t min_particles ) {
  if( g_alt_cap<min_particles ) {
    VPB_CUDA( cudaStreamSynchronize( rt().stream ) );
    if( g_alt ) VPB_CUDA( cudaFree( g_alt ) );
    VPB_CUDA( cudaMalloc( &g_alt, sizeof(vpb_particle_t)*min_particles ) );
    g_alt_cap = min_particles;
  }
  return g_alt;
}

void sort_swap_spare( SpeciesMirror * m ) {
  vpb_particle_t * t = m->d_p; size_t tc = m->cap_p;
  m->d_p = g_alt; m->cap_p = g_alt_cap; g_alt = t; g_alt_cap = tc;
}

void materialize_sort( SpeciesMirror * m ) {
  if( !m->perm_pending ) return;
  Runtime & r = rt();
  m->perm_pending = false;
  if( m->np<=0 ) return;
  vpb_particle_t * out = sort_spare( m->cap_p );
  k_sort_apply_perm<<< div_up( m->np, 256 ), 256, 0, r.stream >>>( m->d_perm, m->np, m->d_p, out );
  VPB_CHECK_LAUNCH(); r.launches++;
  sort_swap_spare( m );
  m->keys_valid = m->keys_patchable = false;   // d_keys is in the old order
}

void release_lazy_sort( SpeciesMirror * m ) {
  if( m->d_perm ) cudaFree( m->d_perm );
  if( m->d_keys ) cudaFree( m->d_keys );
  m->d_perm = m->d_keys = nullptr; m->cap_perm = m->cap_keys = 0; m->perm_pending = m->keys_valid = m->keys_patchable = false;
}

// ---- segment sort: the deterministic stable rank for a nearly sorted array ------------------------------------
//
// Between two sorts the array keeps the order the last sort gave it: position i belongs to the HOME segment of the
// voxel h with home[h] <= i < home[h+1] (home[] = the last sort's partition).  advance_p changes keys, boundary_p
// fills holes and appends, nothing moves a record.  So the particles with key k are, in array order (= the order a
// stable sort must give them):
//     strays whose home is below k  |  the stayers of k's own home segment  |  strays from above and appended ones
// The stayers -- the bulk -- get their final slot from a ballot rank inside their home segment, which one warp
// walks front to back: no atomic, and they arrive in order.  Only the strays take atomic slots in the two small
// regions before and after the stayers, and only those regions (a handful of indices each) are ordered afterwards,
// by one thread per voxel with a sorting network in registers.  (The generic path below -- atomic slots for
// everything, then a warp-wide bitonic network over the whole segment -- spent most of its time ordering
// electron segments that had a few strays among thirty stayers.)
constexpr int SEG_REP = 8;          // consecutive home segments per warp
constexpr int SEG_ROWS = 8;         // rows of 32 keys in flight per warp (eight segments of 32 particles: the warp's whole range)

// A[] = strays from below (before region), B[] = strays from above (after region), C[] = stayers, per key.
template<bool PLACE>
__global__ void __launch_bounds__(256)
k_seg_pass( const int * __restrict__ keys, const int * __restrict__ home, int nv, int np_home,
            int * __restrict__ A, int * __restrict__ B, int * __restrict__ C, int * __restrict__ curb, int * __restrict__ perm ) {
  const int lane = threadIdx.x & 31;
  const int v0 = ( blockIdx.x*( blockDim.x>>5 ) + ( threadIdx.x>>5 ) )*SEG_REP;
  if( v0>=nv ) return;
  // home[v0 .. v0+SEG_REP] clamped to the particles that exist, in lanes 0..SEG_REP
  const int vi = v0+lane<nv ? v0+lane : nv;
  int pl = 0x7fffffff;
  if( lane<=SEG_REP ) { pl = home[vi]; pl = pl<np_home ? pl : np_home; }
  const int b0 = __shfl_sync( 0xffffffffu, pl, 0 ), e0 = __shfl_sync( 0xffffffffu, pl, SEG_REP );
  // COUNT: lane r counts the stayers of segment r.  PLACE: lane r holds the next stayer slot of segment r (A = the
  // stayers' base after the prep pass).
  int mine = 0;
  if( PLACE && lane<SEG_REP && v0+lane<nv ) mine = A[v0+lane];
  const int pl_next = __shfl_down_sync( 0xffffffffu, pl, 1 );          // lane r: where segment r ends
  const unsigned lt = ( 1u<<lane )-1u;
  // (The count pass ran at 91 % issue utilisation with a loop over the segments of every row: everything below is
  //  branch-free per row -- a three-step binary search over the nine boundaries held in lanes 0..8, one ballot.)
  for( int o=b0; o<e0; o+=32*SEG_ROWS ) {
    int k[SEG_ROWS], slot[SEG_ROWS];
    unsigned strays = 0;            // rows in which this lane holds a stray (PLACE: its slot is on its way in slot[])
#   pragma unroll
    for( int u=0; u<SEG_ROWS; u++ ) { const int pos = o + 32*u + lane; k[u] = pos<e0 ? keys[pos] : -1; }
#   pragma unroll
    for( int u=0; u<SEG_ROWS; u++ ) {
      const int ps = o + 32*u, pos = ps + lane;
      slot[u] = 0;
      if( ps>=e0 ) break;
      // home segment of the position: the last boundary <= pos (empty segments are skipped by construction)
      int h = 0, bnd = b0, t;
      t = __shfl_sync( 0xffffffffu, pl, 4 );     if( pos>=t ) { h = 4;  bnd = t; }
      t = __shfl_sync( 0xffffffffu, pl, h+2 );   if( pos>=t ) { h += 2; bnd = t; }
      t = __shfl_sync( 0xffffffffu, pl, h+1 );   if( pos>=t ) { h += 1; bnd = t; }
      const bool valid = k[u]>=0;
      const bool stay = valid && k[u]==v0+h;
      const unsigned sm = __ballot_sync( 0xffffffffu, stay );
      if( PLACE ) {
        // rank among the stayers of the same segment in this row: they sit in the lanes from the segment's first
        // position on (homes are monotone along a row)
        const int first = bnd - ps;
        const unsigned same = first<=0 ? 0xffffffffu : ( 0xffffffffu<<first );
        const int base = __shfl_sync( 0xffffffffu, mine, h );
        if( stay ) perm[ base + __popc( sm & lt & same ) ] = pos;
      }
      {
        // lane r adds the row's stayers of segment r: the lanes [pl - ps, pl_next - ps) of the row
        int lo = pl - ps, hi = pl_next - ps;
        lo = lo<0 ? 0 : lo; hi = hi>32 ? 32 : hi;
        unsigned m = 0;
        if( lane<SEG_REP && hi>lo ) m = ( hi==32 ? 0xffffffffu : ( ( 1u<<hi )-1u ) ) & ~( ( 1u<<lo )-1u );
        mine += __popc( sm & m );
      }
      if( valid && !stay ) {
        const bool below = k[u]>v0+h;                 // its home is below the voxel it is in now
        // (the slots of all rows are requested before the first one is used: one atomic round trip per batch)
        if( PLACE ) { slot[u] = atomicAdd( ( below ? curb : B ) + k[u], 1 ); strays |= 1u<<u; }
        else atomicAdd( ( below ? A : B ) + k[u], 1 );
      }
    }
    if( PLACE ) {
#     pragma unroll
      for( int u=0; u<SEG_ROWS; u++ ) if( strays & ( 1u<<u ) ) perm[ slot[u] ] = o + 32*u + lane;
    }
  }
  if( !PLACE && lane<SEG_REP && v0+lane<nv ) C[v0+lane] = mine;
}

// particles appended since the last sort (positions >= np_home): strays from above for every voxel
template<bool PLACE>
__global__ void __launch_bounds__(256)
k_seg_tail( const int * __restrict__ keys, int first, int np, int * __restrict__ B, int * __restrict__ perm ) {
  const int pos = first + blockIdx.x*blockDim.x + threadIdx.x;
  if( pos>=np ) return;
  const int k = keys[pos];
  if( PLACE ) perm[ atomicAdd( B + k, 1 ) ] = pos; else atomicAdd( B + k, 1 );
}

__global__ void __launch_bounds__(256)
k_seg_sum( const int * __restrict__ A, const int * __restrict__ B, const int * __restrict__ C, int * __restrict__ D, int nv ) {
  const int v = blockIdx.x*blockDim.x + threadIdx.x;
  if( v<nv ) D[v] = A[v] + B[v] + C[v]; else if( v==nv ) D[v] = 0;
}

// after the scan: A = first stayer slot, C = first slot of the after region, B = its cursor
__global__ void __launch_bounds__(256)
k_seg_prep( const int * __restrict__ part, int * __restrict__ A, int * __restrict__ B, int * __restrict__ C, int nv ) {
  const int v = blockIdx.x*blockDim.x + threadIdx.x;
  if( v>=nv ) return;
  const int sb = part[v] + A[v], pa = sb + C[v];
  A[v] = sb; C[v] = pa; B[v] = pa;
}

// ascending order of perm[b .. b+n), n <= N, by one thread: a bitonic network over N registers
template<int N> __device__ __forceinline__ void seg_sort_regs( int * __restrict__ perm, int b, int n ) {
  int x[N];
# pragma unroll
  for( int i=0; i<N; i++ ) x[i] = i<n ? perm[b+i] : 0x7fffffff;
# pragma unroll
  for( int k=2; k<=N; k<<=1 ) {
#   pragma unroll
    for( int j=k>>1; j>0; j>>=1 ) {
#     pragma unroll
      for( int i=0; i<N; i++ ) {
        const int l = i^j;
        if( l>i ) {
          const int lo = min( x[i], x[l] ), hi = max( x[i], x[l] );
          if( ( i & k )==0 ) { x[i] = lo; x[l] = hi; } else { x[i] = hi; x[l] = lo; }
        }
      }
    }
  }
# pragma unroll
  for( int i=0; i<N; i++ ) if( i<n ) perm[b+i] = x[i];
}

constexpr int SEG_SMALL = 16;       // longest region one thread orders

__device__ __forceinline__ void seg_sort_small( int * __restrict__ perm, int b, int n ) {
  if( n==2 ) { const int x0 = perm[b], x1 = perm[b+1]; if( x0>x1 ) { perm[b] = x1; perm[b+1] = x0; } return; }
  if( n<=4 ) { seg_sort_regs<4>( perm, b, n ); return; }
  if( n<=8 ) { seg_sort_regs<8>( perm, b, n ); return; }
  seg_sort_regs<16>( perm, b, n );
}

// ascending order of perm[b .. b+n) by a warp (any n)
__device__ __forceinline__ void seg_sort_warp( int * perm, int b, int n, int lane, int * smem, int * __restrict__ big_scratch ) {
  if( n<=32 )       { fix_gather_regs<1,false>( perm, b, n, lane, nullptr, nullptr ); return; }
  if( n<=64 )       { fix_gather_regs<2,false>( perm, b, n, lane, nullptr, nullptr ); return; }
  if( n<=32*FIX_E ) { fix_gather_regs<FIX_E,false>( perm, b, n, lane, nullptr, nullptr ); return; }
  int n2 = 64; while( n2<n ) n2 <<= 1;
  int * a = n2<=FIX_SMEM ? smem : big_scratch + 2*(size_t)b;
  for( int t=lane; t<n2; t+=32 ) a[t] = t<n ? perm[b+t] : 0x7fffffff;
  __threadfence_block(); __syncwarp();
  bitonic_mem( a, n2, lane );
  for( int t=lane; t<n; t+=32 ) perm[b+t] = a[t];
  __syncwarp();
}

// One thread per voxel orders its two stray regions when they hold up to small_max (<= SEG_SMALL) indices; longer regions
// (a crowded voxel, a test) are ordered by the whole warp, one after the other.
__global__ void __launch_bounds__(32*FIX_WARPS)
k_seg_fix( const int * __restrict__ part, const int * __restrict__ A, const int * __restrict__ C, int nv, int np,
           int * perm, int small_max, int * __restrict__ big_scratch ) {
  __shared__ int s[FIX_WARPS][FIX_SMEM];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int v = blockIdx.x*blockDim.x + threadIdx.x;
  int b[2] = { 0, 0 }, n[2] = { 0, 0 };
  if( v<nv ) {
    const int p0 = part[v], sb = A[v], pa = C[v], p1 = part[v+1];
    b[0] = p0; n[0] = sb-p0; b[1] = pa; n[1] = p1-pa;
  }
# pragma unroll
  for( int q=0; q<2; q++ ) {
    if( n[q]>=2 && n[q]<=small_max ) seg_sort_small( perm, b[q], n[q] );
    unsigned big = __ballot_sync( 0xffffffffu, n[q]>small_max );
    while( big ) {
      const int src = __ffs( big )-1; big &= big-1;
      seg_sort_warp( perm, __shfl_sync( 0xffffffffu, b[q], src ), __shfl_sync( 0xffffffffu, n[q], src ), lane, s[w], big_scratch );
    }
  }
}

static int seg_sort_enabled() {
  static int on = -1;
  if( on<0 ) { const char * e = getenv( "VPIC_B200_SEG_SORT" ); on = ( e && e[0]=='0' ) ? 0 : 1; }
  return on;
}

// regions longer than this are ordered by a warp (VPIC_B200_SEG_SMALL lowers it so that tests reach that path)
static int seg_small_max() {
  static int n = -1;
  if( n<0 ) { const char * e = getenv( "VPIC_B200_SEG_SMALL" ); n = e ? atoi( e ) : SEG_SMALL; if( n<1 ) n = 1; if( n>SEG_SMALL ) n = SEG_SMALL; }
  return n;
}

static int lazy_sort_enabled() {
  static int on = -1;
  if( on<0 ) { const char * e = getenv( "VPIC_B200_LAZY_SORT" ); on = ( e && e[0]=='0' ) ? 0 : 1; }
  return on;
}

} // namespace vpb

using namespace vpb;

// sort_p_pipeline.c:221-371.  Two ways to the same result:
//  * immediate (STRICT coherence, PARITY deposit): histogram of the particles' voxels, scan, index scatter,
//    per-voxel ordering fused with the particle copy into the spare buffer;
//  * lazy (RESIDENT coherence + FAST deposit): the same up to the per-voxel ordering, which writes the ordered
//    INDICES only (d_perm); the particle copy is left to the next advance_p, which reads through d_perm and
//    writes the other buffer while it pushes (k_advance_p<GATHER>), so the sort's 64 B per particle copy pass
//    costs nothing.  If the preceding advance_p left the voxel keys (d_keys), the 32 B records are not read
//    here at all.  Whatever else touches the particles first carries the permutation out (materialize_sort).
extern "C" void sort_p( vpb_species_t * sp ) {
  VPB_TRACE_CALL( "sort_p" );
  if( !sp ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * m = mirror( sp );
  const bool lazy = r.coherence==VPB_COHERENCE_RESIDENT && r.deposit_mode==VPB_DEPOSIT_FAST && lazy_sort_enabled();
  to_device( m );                        // (a sort still pending from an earlier call is carried out here)
  sp->last_sorted = sp->g->step;
  const GridDev & g = m->grid->dev;
  const int np = m->np, nv = g.nv;
  if( !lazy ) sort_spare( m->cap_p );
  else if( np>0 ) {
    // everything the lazy path will ever need is allocated at the species' first sort (start-up), not in the
    // middle of a run: the key array the advance_p before the next sort fills, and the spare particle buffer
    if( m->cap_keys<m->cap_p ) {
      VPB_CUDA( cudaStreamSynchronize( r.stream ) );
      if( m->d_keys ) VPB_CUDA( cudaFree( m->d_keys ) );
      VPB_CUDA( cudaMalloc( &m->d_keys, sizeof(int)*m->cap_p ) ); m->cap_keys = m->cap_p; m->keys_valid = m->keys_patchable = false;
    }
    sort_spare( m->cap_p );
    if( !m->d_partition_alt && seg_sort_enabled() ) VPB_CUDA( cudaMalloc( &m->d_partition_alt, sizeof(int)*( (size_t)nv + 1 ) ) );
  }
  // scratch: keys[np] perm[np] count[nv+1] next[nv+1] tiles big[2np]   (segment sort: five arrays of nv+1, tiles, big)
  const size_t npa = ( (size_t)np + 31 ) & ~(size_t)31, nva = ( (size_t)nv + 1 + 31 ) & ~(size_t)31;
  const size_t ntile = (size_t)div_up( nv+1, SCAN_TILE ) + 32;
  const size_t need_generic = 2*npa + 2*nva + ntile + 2*npa + 64, need_seg = 5*nva + ntile + 2*npa + 64;
  int * base = (int *)scratch( sizeof(int)*( need_generic>need_seg ? need_generic : need_seg ) );
  const bool seg = lazy && np>0 && m->keys_valid && m->d_keys && m->home_valid && seg_sort_enabled();
  if( seg ) {
    // the array still has the order of the last sort and advance_p / boundary_p left the keys: deterministic ranks
    if( m->cap_perm<m->cap_p ) {
      VPB_CUDA( cudaStreamSynchronize( r.stream ) );
      if( m->d_perm ) VPB_CUDA( cudaFree( m->d_perm ) );
      VPB_CUDA( cudaMalloc( &m->d_perm, sizeof(int)*m->cap_p ) ); m->cap_perm = m->cap_p;
    }
    if( !m->d_partition_alt ) VPB_CUDA( cudaMalloc( &m->d_partition_alt, sizeof(int)*( (size_t)nv + 1 ) ) );
    int * A = base, * B = A + nva, * C = B + nva, * D = C + nva, * E = D + nva, * tiles = E + nva, * big = tiles + ntile;
    const int * home = m->d_partition; int * part = m->d_partition_alt;
    const int np_home = m->home_np<np ? m->home_np : np;
    const int pass_blocks = div_up( nv, 8*SEG_REP );
    VPB_CUDA( cudaMemsetAsync( A, 0, sizeof(int)*2*nva, r.stream ) );
    k_seg_pass<false><<< pass_blocks, 256, 0, r.stream >>>( m->d_keys, home, nv, np_home, A, B, C, nullptr, nullptr );
    if( np>np_home ) { k_seg_tail<false><<< div_up( np-np_home, 256 ), 256, 0, r.stream >>>( m->d_keys, np_home, np, B, nullptr ); r.launches++; }
    k_seg_sum<<< div_up( nv+1, 256 ), 256, 0, r.stream >>>( A, B, C, D, nv );
    exclusive_scan( D, part, E, tiles, nv+1 );
    k_seg_prep<<< div_up( nv, 256 ), 256, 0, r.stream >>>( part, A, B, C, nv );
    k_seg_pass<true><<< pass_blocks, 256, 0, r.stream >>>( m->d_keys, home, nv, np_home, A, B, C, E, m->d_perm );
    if( np>np_home ) { k_seg_tail<true><<< div_up( np-np_home, 256 ), 256, 0, r.stream >>>( m->d_keys, np_home, np, B, m->d_perm ); r.launches++; }
    k_seg_fix<<< div_up( nv, 32*FIX_WARPS ), 32*FIX_WARPS, 0, r.stream >>>( part, A, C, nv, np, m->d_perm, seg_small_max(), big );
    VPB_CHECK_LAUNCH(); r.launches += 5;
    m->d_partition_alt = m->d_partition; m->d_partition = part;
    m->perm_pending = true;
    r.sorts_from_keys++; r.seg_sorts++;
  } else {
  int * keys = base, * perm = keys + npa, * count = perm + npa, * next = count + nva, * tiles = next + nva, * big = tiles + ntile;
  VPB_CUDA( cudaMemsetAsync( count, 0, sizeof(int)*( (size_t)nv+1 ), r.stream ) );
  if( np>0 ) {
    if( lazy && m->keys_valid && m->d_keys ) {       // the keys are already there: 4 B per particle instead of 32
      keys = m->d_keys; r.sorts_from_keys++;
      k_hist_keys<<< div_up(np,256*SORT_CHUNKS), 256, 0, r.stream >>>( keys, np, count );
    } else k_sort_hist<<< div_up(np,256*SORT_CHUNKS), 256, 0, r.stream >>>( m->d_p, np, keys, count );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
  exclusive_scan( count, m->d_partition, next, tiles, nv+1 );
  if( np>0 ) {
    if( lazy ) {
      if( m->cap_perm<m->cap_p ) {
        VPB_CUDA( cudaStreamSynchronize( r.stream ) );
        if( m->d_perm ) VPB_CUDA( cudaFree( m->d_perm ) );
        VPB_CUDA( cudaMalloc( &m->d_perm, sizeof(int)*m->cap_p ) ); m->cap_perm = m->cap_p;
      }
      perm = m->d_perm;                // the indices go straight to where advance_p reads them and are put in order in place
    }
    k_sort_scatter_idx<<< div_up(np,256*SORT_CHUNKS), 256, 0, r.stream >>>( keys, np, next, perm );
    const int fix_blocks = div_up( nv, FIX_WARPS*FIX_REP );
    if( lazy ) {
      k_sort_fix_gather<false><<< fix_blocks, 32*FIX_WARPS, 0, r.stream >>>( m->d_partition, nv, np, perm, big, m->d_p, nullptr );
      m->perm_pending = true;
    } else {
      k_sort_fix_gather<true><<< fix_blocks, 32*FIX_WARPS, 0, r.stream >>>( m->d_partition, nv, np, perm, big, m->d_p, g_alt );
      sort_swap_spare( m );            // swap the species buffer with the spare
    }
    VPB_CHECK_LAUNCH(); r.launches += 2;
  }
  }
  m->partition_dev_valid = true;
  const bool keep = m->perm_pending;
  mark_dev_written( m );
  m->perm_pending = keep;
  m->home_valid = true; m->home_np = np;     // d_partition describes the order the array has (once the gather ran)
  if( r.coherence==VPB_COHERENCE_STRICT ) to_host( m );
}

extern "C" long long vpic_b200_sorts_from_keys( void ) { return rt().sorts_from_keys; }
extern "C" long long vpic_b200_segment_sorts( void ) { return rt().seg_sorts; }

// Number of particles that violate "sorted by voxel, partition[] consistent".
// Meaningful right after sort_p (0 expected).  Test / diagnostic entry point.
extern "C" long long vpic_b200_sort_violations( vpb_species_t * sp ) {
  if( !sp ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * m = mirror( sp );
  to_device( m );
  if( !m->partition_dev_valid ) return -1;
  unsigned long long * d_bad = (unsigned long long *)scratch( sizeof(unsigned long long) );
  VPB_CUDA( cudaMemsetAsync( d_bad, 0, sizeof(unsigned long long), r.stream ) );
  if( m->np>0 ) { k_sort_check<<< div_up( m->np, 256 ), 256, 0, r.stream >>>( m->d_p, m->np, m->d_partition, m->grid->dev.nv, d_bad ); VPB_CHECK_LAUNCH(); r.launches++; }
  unsigned long long bad = 0;
  d2h( &bad, d_bad, sizeof(bad), r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  return (long long)bad;
}
