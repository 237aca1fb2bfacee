// vpb_advance_p.cu -- the particle advance: Boris push + charge-conserving
// current deposit + in-kernel move_p.
// Reference: src/species_advance/standard/pipeline/advance_p_pipeline.cc:20-314,
//            src/species_advance/standard/move_p.cc:216-353.
//
// B200 design (FAST deposit), sized from the round-1 ncu capture
// (profiles/r1a_advance_p.md): the first kernel was issue-bound at ~670 thread
// instructions per push, more than half of them spent reducing currents and
// walking divergent move_p loops.  This kernel:
//
//  * A warp owns a TILE of 32*AP_K consecutive particles (8 KB).  One elected
//    lane brings the tile into shared memory with a single bulk asynchronous
//    copy (cp.async.bulk, completion on an mbarrier) and after the push writes
//    the tile back with one bulk store.  No per-particle LDG/STG at all.
//  * Each LANE walks AP_K CONSECUTIVE particles of the tile (its own 256 B run
//    of the array).  With cell-sorted particles a lane stays in one voxel for
//    most of its run, so it keeps that voxel's 18 interpolator coefficients
//    and the 12 running current sums in REGISTERS, and only when the voxel
//    changes commits the sums with three 128-bit vector reductions
//    (red.global.add.v4.f32) and re-gathers the coefficients.  Deposit cost
//    drops from ~250 to ~15 instructions per push.
//  * Particles that leave their voxel are not moved in line (that loop is
//    divergent and ran at <10 % lane utilisation).  They are pushed on a
//    per-warp queue in shared memory; whenever 32 are waiting the whole warp
//    runs move_p on them with every lane busy.
//
// Algorithmic HBM traffic: 32 B read + 32 B written per particle plus, per
// voxel, 80 B interpolator + 48 B accumulator read-modify-write (SURVEY 8d:
// 64 + 176/ppc bytes per push).
#include "vpb_push.cuh"
#include <vector>
#include <cstdlib>

namespace vpb {

#ifndef VPB_AP_K
#define VPB_AP_K 8
#endif
#ifndef VPB_AP_NBUF
#define VPB_AP_NBUF 1
#endif
#ifndef VPB_AP_MINBLK
#define VPB_AP_MINBLK 10
#endif
constexpr int AP_K      = VPB_AP_K;     // consecutive particles per lane
constexpr int AP_NBUF   = VPB_AP_NBUF;  // tile buffers per warp: 2 = the next tile is loading while this one is pushed
constexpr int AP_TILE   = 32*AP_K;      // particles per warp tile
#ifndef VPB_AP_WARPS
#define VPB_AP_WARPS 2
#endif
constexpr int AP_WARPS  = VPB_AP_WARPS;            // warps per block
constexpr int AP_MINBLK = VPB_AP_MINBLK;            // resident blocks per SM the registers are budgeted for
constexpr int AP_QCAP   = 64;           // crosser queue entries per warp

struct __align__(128) WarpTile {
  float4 tile[AP_NBUF][2*AP_TILE];      // the particles, array order (8 KB per buffer)
  // queue of particles that left their voxel: remaining displacement + index
  // (position, voxel, momentum and weight are still in the particle's slot)
  float4 q[AP_QCAP];                    // dispx, dispy, dispz, index relative to the launch's first particle
  unsigned long long bar[AP_NBUF];      // mbarrier of each buffer's tile load
};

__device__ __forceinline__ uint32_t smem_u32( const void * p ) { return (uint32_t)__cvta_generic_to_shared( p ); }

// Run move_p (move_p.cc:216-353) on up to 32 queued particles, one per lane.
// Entries of the tile that is still in shared memory are finished there;
// entries carried over from tiles already written back are finished in global
// memory (the caller has waited for those bulk stores to complete).
// `on_chip( rel )` returns the particle's slot in shared memory, or nullptr.
template< class OnChip >
__device__ __forceinline__ void drain_crossers( const float4 * __restrict__ queue, OnChip on_chip, int base, int n, int lane,
                                                vpb_particle_t * __restrict__ p, int idx_base,
                                                float * __restrict__ acc, vpb_particle_mover_t * __restrict__ pm,
                                                int max_nm, int * __restrict__ counters, const GridDev & g, float qsp,
                                                int * __restrict__ keys = nullptr ) {
  if( lane>=n ) return;
  const int e = base + lane;
  const float4 qe = queue[e];
  const int rel = __float_as_int( qe.w );
  float4 * slot = on_chip( rel );
  const bool off_chip = !slot;
  if( !slot ) slot = reinterpret_cast<float4 *>( p + rel );
  const float4 r0 = slot[0], r1 = slot[1];
  MoveState ms;
  ms.dx = r0.x; ms.dy = r0.y; ms.dz = r0.z; ms.i = __float_as_int( r0.w );
  ms.ux = r1.x; ms.uy = r1.y; ms.uz = r1.z;
  ms.dispx = qe.x; ms.dispy = qe.y; ms.dispz = qe.z;
  const float q = r1.w*qsp;
  float j[12];
  int status;
  do {
    int vox;
    status = move_p_streak( ms, q, g, j, vox );
    commit_currents( acc, vox, j );
  } while( status==1 );
  slot[0] = make_float4( ms.dx, ms.dy, ms.dz, __int_as_float( ms.i ) );
  slot[1] = make_float4( ms.ux, ms.uy, ms.uz, r1.w );       // reflections flip a momentum component
  if( keys && off_chip ) keys[rel] = ms.i;                  // its tile's keys were written with the tile
  if( status==2 ) {
    const int k = atomicAdd( counters+0, 1 );
    if( k<max_nm ) reinterpret_cast<float4 *>( pm )[k] = make_float4( ms.dispx, ms.dispy, ms.dispz, __int_as_float( idx_base + rel ) );
    else atomicAdd( counters+1, 1 );
  }
}

#ifndef VPB_AP_DRAIN_MIN
#define VPB_AP_DRAIN_MIN 24
#endif
constexpr int AP_DRAIN_MIN = VPB_AP_DRAIN_MIN;        // fewer crossers than this wait for the next tile's

// p points at the first particle of the range this launch owns (np particles,
// ntile tiles); idx_base is that particle's index in sp->p (mover indices).
//
// GATHER: the launch that follows a lazy sort_p (vpb_sort.cu).  sort_p has only
// computed the sorted order, perm[d] = index of the particle that belongs in
// slot d; the tile for slots [t0, t0+cnt) is gathered from src[perm[..]] with
// 256-bit loads instead of one bulk copy, pushed, and bulk-stored to p[t0..]
// (the species' other buffer).  The sort's own 64 B per particle copy pass
// disappears into the push's read and write.
// KEYS: the launch before a sort_p also leaves every particle's final voxel
// index in keys[] (4 B per particle), so that sort_p does not have to touch
// the 32 B particle records to learn the keys.
template< bool GATHER, bool KEYS >
__global__ void __launch_bounds__( 32*AP_WARPS, AP_MINBLK )
k_advance_p( vpb_particle_t * __restrict__ p, int np, int ntile, int idx_base,
             const float * __restrict__ ip, float * __restrict__ acc,
             vpb_particle_mover_t * __restrict__ pm, int max_nm, int * __restrict__ counters,
             GridDev g, PushCoef c,
             const vpb_particle_t * __restrict__ src, const int * __restrict__ perm, int * __restrict__ keys ) {
  __shared__ WarpTile s_all[AP_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpTile & S = s_all[warp];
  const int total_warps = gridDim.x*AP_WARPS;
  const int gw = blockIdx.x*AP_WARPS + warp;
  if( gw>=ntile ) return;
  if( lane==0 ) {
#   pragma unroll
    for( int b=0; b<AP_NBUF; b++ ) asm volatile( "mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32( &S.bar[b] )) : "memory" );
    asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
    asm volatile( "fence.proxy.async.shared::cta;" ::: "memory" );
  }
  __syncwarp();

  // one elected lane starts the bulk copy of a tile into buffer b
  auto start_load = [&]( int tile, int b ) {
    const long long t0 = (long long)tile*AP_TILE;
    const uint32_t bytes = (uint32_t)( ( np - t0 < AP_TILE ) ? (int)( np - t0 ) : AP_TILE )*32u;
    const uint32_t bar = smem_u32( &S.bar[b] ), dst = smem_u32( S.tile[b] );
    asm volatile( "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory" );
    asm volatile( "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                  :: "r"(dst), "l"(p + t0), "r"(bytes), "r"(bar) : "memory" );
  };

  uint32_t phase = 0;                          // bit b: parity the next wait on buffer b expects
  int buf = 0;
  int qn = 0;                                  // queued crossers (warp uniform)
  int n_carry = 0;                             // of which: from tiles whose bulk store may still be in flight
  int cur = -1;                                // voxel whose current sums this lane holds in run[]
  int fvox = -1;                               // voxel whose coefficients this lane holds in f
  // run[] is never zeroed in a branch (that made the compiler copy all twelve
  // sums twice per particle): a new run starts with keepf = 0 and its first
  // deposit is run*0 + j; keepf = 1 afterwards makes the FMA an exact add.
  // keepf != 0 also says "run[] holds sums that are not committed yet".
  float keepf = 0.f;
  float run[12];
# pragma unroll
  for( int k=0; k<12; k++ ) run[k] = 0.f;
  Interp f = {};

  // GATHER: the source indices of the tile's slots, requested one tile ahead (the loads of the NEXT tile's
  // indices are issued when this tile's pushes are done and land while it is stored)
  int sidx[GATHER ? AP_K : 1];
  auto load_indices = [&]( int tile ) {
    const long long t0 = (long long)tile*AP_TILE;
#   pragma unroll
    for( int i=0; i<( GATHER ? AP_K : 1 ); i++ ) { const long long d = t0 + 32*i + lane; sidx[i] = ( tile<ntile && d<np ) ? __ldg( perm + d ) : -1; }
  };
  if( GATHER ) load_indices( gw );
  if( AP_NBUF>1 && lane==0 ) start_load( gw, 0 );
  for( int tile=gw; tile<ntile; tile+=total_warps ) {
    const long long t0 = (long long)tile*AP_TILE;
    const int cnt = ( np - t0 < AP_TILE ) ? (int)( np - t0 ) : AP_TILE;
    if( !GATHER && AP_NBUF==1 && lane==0 ) start_load( tile, 0 );
    float4 * __restrict__ T = S.tile[buf];
    auto on_chip = [&]( int rel ) -> float4 * { return ( rel>=t0 && rel<t0+cnt ) ? &T[ 2*(int)( rel-t0 ) ] : nullptr; };
    if( GATHER ) {
      // 32 consecutive slots per step: coalesced index loads, then every particle record goes from global to
      // shared memory as two 16-byte asynchronous copies (cp.async: no registers, no wait between them -- with
      // plain loads each record's shared store waited for its own load, eight exposed latencies per tile, and
      // the launch was 26 % slower than a push of the same particles in place); consecutive lanes fill
      // consecutive 32 B of the tile, and read consecutive records for the particles that did not move since
      // the last sort.
#     pragma unroll
      for( int i=0; i<AP_K; i++ ) {
        const int d = 32*i + lane;
        if( sidx[i]>=0 ) {
          const uint32_t dst = smem_u32( &T[2*d] );
          const char * s8 = reinterpret_cast<const char *>( src + sidx[i] );
          asm volatile( "cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst),      "l"(s8)      : "memory" );
          asm volatile( "cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst+16u),  "l"(s8+16)   : "memory" );
        }
      }
      asm volatile( "cp.async.commit_group;" ::: "memory" );
      asm volatile( "cp.async.wait_group 0;" ::: "memory" );
      __syncwarp();
    } else {
      const uint32_t bar = smem_u32( &S.bar[buf] );
      uint32_t ok;
      do {
        asm volatile( "{ .reg .pred P1; mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2; selp.u32 %0, 1, 0, P1; }"
                      : "=r"(ok) : "r"(bar), "r"( ( phase>>buf )&1u ) : "memory" );
      } while( !ok );
      phase ^= 1u<<buf;
    }

#   pragma unroll 1
    for( int k=0; k<AP_K; k++ ) {
      if( AP_NBUF>1 && k==AP_K/2 && tile+total_warps<ntile ) {
        // the other buffer's store was issued half a tile ago: its shared-memory
        // read is long done, so the next tile can start loading into it now
        if( lane==0 ) {
          asm volatile( "cp.async.bulk.wait_group.read 0;" ::: "memory" );
          start_load( tile+total_warps, buf^1 );
        }
      }
      // rotate the start of each lane's run so the 128-bit shared accesses of a
      // quarter warp spread over the banks
      const int loc = lane*AP_K + ( ( k + lane ) & ( AP_K-1 ) );
      bool cross = false;
      float cux = 0, cuy = 0, cuz = 0;
      const int qtag = (int)t0 + loc;
      if( loc<cnt ) {
        const float4 r0 = T[2*loc], r1 = T[2*loc+1];
        const float dx = r0.x, dy = r0.y, dz = r0.z; const int ii = __float_as_int( r0.w );
        float ux = r1.x, uy = r1.y, uz = r1.z;
        if( ii!=fvox ) { f = load_interp( ip, ii ); fvox = ii; }
        if( ii!=cur ) {
          // A particle that strayed into this run since the last sort_p (the
          // lane's next particle is back in the run's voxel) deposits on its
          // own below and leaves the running sums alone.
          bool stray = false;
          if( k+1<AP_K ) {
            const int nloc = lane*AP_K + ( ( k + 1 + lane ) & ( AP_K-1 ) );
            stray = nloc<cnt && __float_as_int( T[2*nloc].w )==cur;
          }
          if( !stray ) {
            if( keepf!=0.f ) commit_currents( acc, cur, run );
            cur = ii; keepf = 0.f;
          }
        }
        const float one = 1.0f, one_third = 1.0/3.0;
        float v0, x1, b2;
        boris_inrange( f, c, dx, dy, dz, ux, uy, uz, x1, b2 );
        const float x2 = one + ( ux*ux + ( uy*uy + uz*uz ) );
        v0 = rcp_inrange( sqrt_inrange( x2 ) );
        if( __builtin_expect( !( x1<=VPB_INRANGE_MAX ) || !( b2<=VPB_INRANGE_MAX ) || !( x2<=VPB_INRANGE_MAX ) || c.guarded, 0 ) ) {
          ux = r1.x; uy = r1.y; uz = r1.z;                   // out of range (or NaN): the guarded operators
          boris( f, c, dx, dy, dz, ux, uy, uz );
          v0 = one/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
        }
        const float sux = ux, suy = uy, suz = uz;            // momentum to store
        ux *= c.cdt_dx; uy *= c.cdt_dy; uz *= c.cdt_dz;
        ux *= v0; uy *= v0; uz *= v0;
        v0 = dx + ux; const float v1 = dy + uy, v2 = dz + uz;  // streak midpoint
        const float v3 = v0 + ux, v4 = v1 + uy, v5 = v2 + uz;  // new position
        const float q = r1.w*c.qsp;
        T[2*loc+1] = make_float4( sux, suy, suz, r1.w );
        if( fmaxf( fmaxf( fabsf( v3 ), fabsf( v4 ) ), fabsf( v5 ) )<=one ) {   // all six faces at once (move_p otherwise)
          T[2*loc] = make_float4( v3, v4, v5, r0.w );
          float j[12];
          streak_currents_fast( j, q, v0, v1, v2, ux, uy, uz, q*ux*uy*uz*one_third );
          if( ii==cur ) {
#           pragma unroll
            for( int m=0; m<12; m++ ) run[m] = __fmaf_rn( run[m], keepf, j[m] );
            keepf = 1.f;
          } else commit_currents( acc, ii, j );
        } else {
          cross = true;
          cux = ux; cuy = uy; cuz = uz;
        }
      }
      const unsigned m = __ballot_sync( 0xffffffffu, cross );
      if( m ) {
        if( cross ) {
          const int e = qn + __popc( m & ( ( 1u<<lane )-1u ) );
          S.q[e] = make_float4( cux, cuy, cuz, __int_as_float( qtag ) );
        }
        qn += __popc( m );
        __syncwarp();
        if( qn>=32 ) {
          qn -= 32;
          if( qn<n_carry ) {       // this batch reaches entries of tiles already written back: their stores must have landed
            if( lane==0 ) asm volatile( "cp.async.bulk.wait_group 0;" ::: "memory" );
            __syncwarp();
            n_carry = qn;
          }
          drain_crossers( S.q, on_chip, qn, 32, lane, p, idx_base, acc, pm, max_nm, counters, g, c.qsp, KEYS ? keys : nullptr );
          __syncwarp();
        }
      }
    }

    // finish the queued crossers while the tile is still on chip -- unless there are only a few
    if( qn>=AP_DRAIN_MIN ) {
      if( n_carry>0 ) {
        if( lane==0 ) asm volatile( "cp.async.bulk.wait_group 0;" ::: "memory" );
        __syncwarp();
        n_carry = 0;
      }
      while( qn>0 ) {
        const int n = qn<32 ? qn : 32;
        qn -= n;
        drain_crossers( S.q, on_chip, qn, n, lane, p, idx_base, acc, pm, max_nm, counters, g, c.qsp, KEYS ? keys : nullptr );
        __syncwarp();
      }
    }
    if( GATHER ) load_indices( tile+total_warps );
    if( KEYS ) {
      // the voxel index is the 4th word of every record of the tile: 32 consecutive keys per store
      // (particles that left the domain carry 8*voxel+face and are boundary_p's, which invalidates the keys)
      __syncwarp();
#     pragma unroll
      for( int i=0; i<AP_K; i++ ) { const int d = 32*i + lane; if( d<cnt ) keys[t0+d] = __float_as_int( T[2*d].w ); }
    }
    n_carry = qn;              // whatever is left rides along; its tile is written back now
    // write the tile back: generic-proxy writes -> async-proxy bulk store
    asm volatile( "fence.proxy.async.shared::cta;" ::: "memory" );
    __syncwarp();
    if( lane==0 ) {
      const uint32_t bytes = (uint32_t)cnt*32u;
      asm volatile( "cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(p + t0), "r"(smem_u32( T )), "r"(bytes) : "memory" );
      asm volatile( "cp.async.bulk.commit_group;" ::: "memory" );
      if( AP_NBUF==1 ) asm volatile( "cp.async.bulk.wait_group.read 0;" ::: "memory" );   // the tile buffer is reused at once
    }
    __syncwarp();
    if( AP_NBUF>1 ) buf ^= 1;
  }

  if( qn>0 ) {                 // all tiles are stored: finish the stragglers in global memory
    if( lane==0 ) asm volatile( "cp.async.bulk.wait_group 0;" ::: "memory" );
    __syncwarp();
    drain_crossers( S.q, []( int ) -> float4 * { return nullptr; }, 0, qn, lane, p, idx_base, acc, pm, max_nm, counters, g, c.qsp, KEYS ? keys : nullptr );
  }
  if( keepf!=0.f ) commit_currents( acc, cur, run );
}

} // namespace vpb

using namespace vpb;

namespace {

template< bool GATHER, bool KEYS >
int advance_p_blocks_per_sm() {
  static int blocks_per_sm = 0;
  if( !blocks_per_sm ) {
    auto kern = k_advance_p<GATHER,KEYS>;
    VPB_CUDA( cudaFuncSetAttribute( kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared ) );
    VPB_CUDA( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &blocks_per_sm, kern, 32*AP_WARPS, 0 ) );
    if( blocks_per_sm<1 ) VPB_ERROR( "k_advance_p does not fit on an SM" );
    // Ask for no more shared memory than the resident blocks use (tiles + 1 KB
    // per block that the driver reserves): what is left of the 256 KB is L1
    // for the interpolator gathers.  VPIC_B200_AP_CARVEOUT (percent) overrides.
    int pct = (int)( ( (size_t)blocks_per_sm*( sizeof(WarpTile)*AP_WARPS + 1024 )*100 + 228*1024 - 1 )/( 228*1024 ) );
    if( const char * e = getenv( "VPIC_B200_AP_CARVEOUT" ) ) pct = atoi( e );
    if( pct>100 ) pct = 100;
    VPB_CUDA( cudaFuncSetAttribute( kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct ) );
    int check = 0;
    VPB_CUDA( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &check, kern, 32*AP_WARPS, 0 ) );
    if( check<blocks_per_sm ) VPB_CUDA( cudaFuncSetAttribute( kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared ) );
  }
  return blocks_per_sm;
}

// dst: the buffer the pushed particles are written to (= the buffer they are read from unless src / perm
// say otherwise: the launch after a lazy sort_p); keys: receives the voxel of every particle, or nullptr
template< bool GATHER, bool KEYS >
void launch_advance_p_t( SpeciesMirror * ms, AccMirror * ma, InterpMirror * mi, const PushCoef & c, vpb_particle_t * dst, int first, int count,
                         const vpb_particle_t * src, const int * perm, int * keys ) {
  Runtime & r = rt();
  const int ntile = div_up( count, AP_TILE );
  int nblock = r.sm_count*advance_p_blocks_per_sm<GATHER,KEYS>();
  if( nblock>div_up( ntile, AP_WARPS ) ) nblock = div_up( ntile, AP_WARPS );
  k_advance_p<GATHER,KEYS><<< nblock, 32*AP_WARPS, 0, r.stream >>>( dst + first, count, ntile, first, mi->d, ma->d, ms->d_pm, (int)ms->cap_pm,
                                                                     ms->d_counters, ms->grid->dev, c, src, perm, keys );
  VPB_CHECK_LAUNCH(); r.launches++;
}

void launch_advance_p( SpeciesMirror * ms, AccMirror * ma, InterpMirror * mi, const PushCoef & c, int first, int count, int * keys = nullptr ) {
  if( keys ) launch_advance_p_t<false,true >( ms, ma, mi, c, ms->d_p, first, count, nullptr, nullptr, keys );
  else       launch_advance_p_t<false,false>( ms, ma, mi, c, ms->d_p, first, count, nullptr, nullptr, nullptr );
}

// STRICT coherence with host-resident particles: stream the array through the
// device in chunks -- upload chunk k+1, push chunk k, download chunk k-1 run
// concurrently on three streams (PCIe is full duplex), so the call costs about
// max(upload, download) instead of upload + push + download.
constexpr int AP_CHUNK = 1<<22;       // particles per chunk (128 MB)

bool advance_p_streamed( SpeciesMirror * ms, AccMirror * ma, InterpMirror * mi, const PushCoef & c ) {
  Runtime & r = rt();
  vpb_species_t * sp = ms->host;
  if( r.coherence!=VPB_COHERENCE_STRICT || !sp->p || sp->np<2*AP_CHUNK || r.deposit_mode!=VPB_DEPOSIT_FAST ) return false;
  if( ms->dev_valid && !ms->host_valid ) return false;             // the device holds the only current copy
  // TRACKED coherence: the host program has not written the particles since the library last saw them, so the
  // device copy is current and only the download half of the pipeline runs
  const bool clean = ms->dev_valid && ms->np==sp->np && ms->nm==sp->nm && track_clean( sp->p, sizeof(vpb_particle_t)*(size_t)sp->np );
  if( clean ) materialize_sort( ms );
  ms->np = sp->np;
  pin_host( sp->p, sizeof(vpb_particle_t)*(size_t)sp->max_np );
  static std::vector<cudaEvent_t> ev;
  const int nchunk = div_up( ms->np, AP_CHUNK );
  while( (int)ev.size()<2*nchunk+1 ) { cudaEvent_t e; VPB_CUDA( cudaEventCreateWithFlags( &e, cudaEventDisableTiming ) ); ev.push_back( e ); }
  VPB_CUDA( cudaEventRecord( ev[2*nchunk], r.stream ) );            // uploads must not overtake earlier work on the buffers
  VPB_CUDA( cudaStreamWaitEvent( r.copy_stream[0], ev[2*nchunk], 0 ) );
  for( int k=0; k<nchunk; k++ ) {
    const int first = k*AP_CHUNK, count = ( ms->np-first<AP_CHUNK ) ? ms->np-first : AP_CHUNK;
    const size_t bytes = sizeof(vpb_particle_t)*(size_t)count;
    if( !clean ) {
      h2d( ms->d_p + first, sp->p + first, bytes, r.copy_stream[0] );
      VPB_CUDA( cudaEventRecord( ev[2*k], r.copy_stream[0] ) );
      VPB_CUDA( cudaStreamWaitEvent( r.stream, ev[2*k], 0 ) );
    }
    launch_advance_p( ms, ma, mi, c, first, count );
    VPB_CUDA( cudaEventRecord( ev[2*k+1], r.stream ) );
    VPB_CUDA( cudaStreamWaitEvent( r.copy_stream[1], ev[2*k+1], 0 ) );
    d2h( sp->p + first, ms->d_p + first, bytes, r.copy_stream[1] );
  }
  VPB_CUDA( cudaStreamSynchronize( r.copy_stream[1] ) );
  ms->dev_valid = true; ms->host_valid = true; ms->partition_dev_valid = false;
  ms->perm_pending = false; ms->keys_valid = ms->keys_patchable = false; ms->home_valid = false;
  track_arm( sp->p, sizeof(vpb_particle_t)*(size_t)sp->max_np );
  return true;
}

} // namespace

extern "C" void advance_p( vpb_species_t * sp, vpb_accumulator_array_t * aa, const vpb_interpolator_array_t * ia ) {
  VPB_TRACE_CALL( "advance_p" );
  if( !sp || !aa || !ia || sp->g!=aa->g || sp->g!=ia->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( sp );
  AccMirror     * ma = mirror( aa );
  InterpMirror  * mi = mirror( const_cast<vpb_interpolator_array_t *>( ia ) );
  particle_grid( ms->grid );
  to_device( ma ); to_device( mi );
  const vpb_grid_t * hg = sp->g;
  PushCoef c;
  c.qdt_2mc = (sp->q*hg->dt)/(2*sp->m*hg->cvac);      // advance_p_pipeline.cc:267-271
  c.cdt_dx  = hg->cvac*hg->dt*hg->rdx;
  c.cdt_dy  = hg->cvac*hg->dt*hg->rdy;
  c.cdt_dz  = hg->cvac*hg->dt*hg->rdz;
  c.qsp     = sp->q;
  { const float a = fabsf( c.qdt_2mc ); c.guarded = ( a>=9.31322574615478515625e-10f && a<=1073741824.0f ) ? 0 : 1; }
  VPB_CUDA( cudaMemsetAsync( ms->d_counters, 0, sizeof(int)*2, r.stream ) );
  const bool streamed = advance_p_streamed( ms, ma, mi, c );
  bool keys_written = false;
  if( !streamed ) {
    const bool fast = r.deposit_mode==VPB_DEPOSIT_FAST;
    to_device( ms, /*keep_lazy_sort=*/fast );
    if( ms->np>0 ) {
      if( !fast ) advance_p_parity( ms, ma, mi, c );
      else {
        // The launch before a sort_p (advance.cc:25-29 sorts at the top of the step whose number is a multiple of
        // sort_interval) also writes the voxel keys, so that a lazy sort_p need not read the particle records.
        int * keys = nullptr;
        if( r.coherence==VPB_COHERENCE_RESIDENT && sp->sort_interval>0 && ( ( hg->step+1 ) % sp->sort_interval )==0 ) {
          if( ms->cap_keys<ms->cap_p ) {
            VPB_CUDA( cudaStreamSynchronize( r.stream ) );
            if( ms->d_keys ) VPB_CUDA( cudaFree( ms->d_keys ) );
            VPB_CUDA( cudaMalloc( &ms->d_keys, sizeof(int)*ms->cap_p ) ); ms->cap_keys = ms->cap_p;
          }
          keys = ms->d_keys;
        }
        if( ms->perm_pending ) {          // sort_p left the order in d_perm: gather through it, write the other buffer
          vpb_particle_t * dst = sort_spare( ms->cap_p );
          if( keys ) launch_advance_p_t<true,true >( ms, ma, mi, c, dst, 0, ms->np, ms->d_p, ms->d_perm, keys );
          else       launch_advance_p_t<true,false>( ms, ma, mi, c, dst, 0, ms->np, ms->d_p, ms->d_perm, nullptr );
          sort_swap_spare( ms );
          ms->perm_pending = false;
        } else launch_advance_p( ms, ma, mi, c, 0, ms->np, keys );
        keys_written = keys!=nullptr;
      }
    }
  }
  // The mover count is what the host driver reads next (advance.cc:73-101).
  int cnt[2];
  d2h( cnt, ms->d_counters, sizeof(int)*2, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( cnt[1] ) VPB_WARNING( "Species \"%s\" ran out of storage for %i movers", sp->name, cnt[1] );   // advance_p_pipeline.cc:301-305
  ms->nm = cnt[0]<(int)ms->cap_pm ? cnt[0] : (int)ms->cap_pm;
  if( ms->nm>1 ) sort_movers_by_index( ms );
  ms->movers_sorted = true;
  const bool home = ms->home_valid;          // a push moves no record: the array keeps the order of the last sort
  if( !streamed ) { mark_dev_written( ms ); ms->home_valid = home; }
  // Particles on the mover list carry 8*voxel+face: their keys are wrong until boundary_p has removed them, which
  // patches d_keys at the holes it fills and the records it appends (k_keys_patch).  A mover that did not fit the
  // list (cnt[1]) is in that state for good.
  ms->keys_valid     = keys_written && ms->nm==0 && !cnt[1];
  ms->keys_patchable = keys_written && ms->nm>0  && !cnt[1];
  mark_dev_written( ma );
  movers_to_host( ms );
  if( r.coherence==VPB_COHERENCE_STRICT ) { to_host( ms ); to_host( ma ); }
}
