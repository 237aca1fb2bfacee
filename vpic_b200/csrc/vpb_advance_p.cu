// vpb_advance_p.cu -- the particle advance: Boris push + charge-conserving
// current deposit + in-kernel move_p.
// Reference: src/species_advance/standard/pipeline/advance_p_pipeline.cc:20-314,
//            src/species_advance/standard/move_p.cc:216-353.
//
// B200 design (FAST deposit):
//  * particle_t is streamed as two 128-bit loads / stores per particle; a warp
//    owns a STRIP of consecutive particles split into two half strips (lanes
//    0-15 / 16-31) so every half-warp access is a contiguous 512 B run.
//  * the 18 interpolator coefficients are gathered with read-only 128-bit loads;
//    with cell-sorted particles neighbouring lanes hit the same 80 B record so
//    the gather is an L1 broadcast.
//  * deposit: the 12 currents of the 32 particles of one iteration are staged in
//    shared memory; 24 "walker" lanes (12 components x 2 half strips) then add
//    them up sequentially along the strip, carrying the running sum across
//    iterations while the voxel stays the same, and commit one red.global.add
//    per component only when the voxel changes.  For sorted particles that is
//    ~12 atomics per (voxel, warp) instead of 12 per particle.
//  * particles that leave their voxel run move_p in the kernel (divergent, rare)
//    and deposit their streaks with plain atomics; those that leave the rank's
//    domain are appended to the mover list for boundary_p.
// Algorithmic HBM traffic: 32 B read + 32 B written per particle plus, per
// voxel, 80 B interpolator + 48 B accumulator read-modify-write (SURVEY 8d:
// 64 + 176/ppc bytes per push).
#include "vpb_push.cuh"

namespace vpb {

constexpr int AP_WARPS   = 8;                 // warps per block
constexpr int AP_ITERS   = 8;                 // iterations per strip
constexpr int AP_STRIP   = 32*AP_ITERS;       // particles per warp
constexpr int AP_HALF    = 16*AP_ITERS;       // particles per half strip
constexpr int AP_SSTRIDE = 33;                // smem row stride (bank-conflict free)

__global__ void __launch_bounds__( 32*AP_WARPS )
k_advance_p( vpb_particle_t * __restrict__ p, int np,
             const float * __restrict__ ip, float * __restrict__ acc,
             vpb_particle_mover_t * __restrict__ pm, int max_nm, int * __restrict__ counters,
             GridDev g, PushCoef c ) {
  __shared__ float s_val[AP_WARPS][12*AP_SSTRIDE];
  __shared__ int   s_vox[AP_WARPS][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int half = lane >> 4, hl = lane & 15;
  const long long strip0 = ( (long long)blockIdx.x*AP_WARPS + warp )*AP_STRIP;
  if( strip0>=np ) return;
  float * sv = s_val[warp]; int * sx = s_vox[warp];
  // walker state: lane (half, comp) sums component comp of its half strip
  const int comp = hl;                        // 0..11 active
  int cur = -1; float run = 0.f;

  for( int it=0; it<AP_ITERS; it++ ) {
    const long long idx = strip0 + (long long)half*AP_HALF + it*16 + hl;
    const bool valid = idx<np;
    float j[12];
    int dep_vox = -1;
    bool crosser = false;
    MoveState ms; float q = 0.f, w_keep = 0.f;
    if( valid ) {
      float4 * pp = reinterpret_cast<float4 *>( p + idx );
      const float4 r0 = pp[0], r1 = pp[1];
      float dx = r0.x, dy = r0.y, dz = r0.z; const int ii = __float_as_int( r0.w );
      float ux = r1.x, uy = r1.y, uz = r1.z; q = r1.w; w_keep = r1.w;
      const Interp f = load_interp( ip, ii );
      boris( f, c, dx, dy, dz, ux, uy, uz );
      const float sux = ux, suy = uy, suz = uz;            // momentum to store
      const float one = 1.0f, one_third = 1.0/3.0;
      float v0 = one/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
      ux *= c.cdt_dx; uy *= c.cdt_dy; uz *= c.cdt_dz;
      ux *= v0; uy *= v0; uz *= v0;
      v0 = dx + ux; const float v1 = dy + uy, v2 = dz + uz;  // streak midpoint
      const float v3 = v0 + ux, v4 = v1 + uy, v5 = v2 + uz;  // new position
      q *= c.qsp;
      if( v3<=one && v4<=one && v5<=one && -v3<=one && -v4<=one && -v5<=one ) {
        pp[0] = make_float4( v3, v4, v5, r0.w );
        pp[1] = make_float4( sux, suy, suz, r1.w );
        streak_currents( j, q, v0, v1, v2, ux, uy, uz, q*ux*uy*uz*one_third );
        dep_vox = ii;
      } else {
        crosser = true;
        ms.dx = dx; ms.dy = dy; ms.dz = dz; ms.i = ii;
        ms.ux = sux; ms.uy = suy; ms.uz = suz;
        ms.dispx = ux; ms.dispy = uy; ms.dispz = uz;
      }
    }
    // ---- stage the in-voxel currents and walk them ----
    sx[lane] = dep_vox;
    if( dep_vox>=0 ) {
#     pragma unroll
      for( int k=0; k<12; k++ ) sv[ k*AP_SSTRIDE + lane ] = j[k];
    }
    __syncwarp();
    if( comp<12 ) {
      const int * vx = sx + 16*half;
      const float * vv = sv + comp*AP_SSTRIDE + 16*half;
#     pragma unroll
      for( int k=0; k<16; k++ ) {
        const int vox = vx[k];
        if( vox>=0 ) {
          if( vox!=cur ) {
            if( cur>=0 ) atomicAdd( acc + (size_t)cur*ACC_STRIDE + comp, run );
            cur = vox; run = 0.f;
          }
          run += vv[k];
        }
      }
    }
    __syncwarp();
    // ---- particles that left their voxel: move_p, streak by streak ----
    if( crosser ) {
      int status;
      do {
        int vox;
        status = move_p_streak( ms, q, g, j, vox );
        float * a = acc + (size_t)vox*ACC_STRIDE;
#       pragma unroll
        for( int k=0; k<12; k++ ) atomicAdd( a+k, j[k] );
      } while( status==1 );
      float4 * pp = reinterpret_cast<float4 *>( p + idx );
      pp[0] = make_float4( ms.dx, ms.dy, ms.dz, __int_as_float( ms.i ) );
      pp[1] = make_float4( ms.ux, ms.uy, ms.uz, w_keep );
      if( status==2 ) {
        const int slot = atomicAdd( counters+0, 1 );
        if( slot<max_nm ) {
          reinterpret_cast<float4 *>( pm )[slot] = make_float4( ms.dispx, ms.dispy, ms.dispz, __int_as_float( (int)idx ) );
        } else atomicAdd( counters+1, 1 );
      }
    }
  }
  if( comp<12 && cur>=0 ) atomicAdd( acc + (size_t)cur*ACC_STRIDE + comp, run );
}

} // namespace vpb

using namespace vpb;

extern "C" void advance_p( vpb_species_t * sp, vpb_accumulator_array_t * aa, const vpb_interpolator_array_t * ia ) {
  if( !sp || !aa || !ia || sp->g!=aa->g || sp->g!=ia->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( sp );
  AccMirror     * ma = mirror( aa );
  InterpMirror  * mi = mirror( const_cast<vpb_interpolator_array_t *>( ia ) );
  to_device( ms ); to_device( ma ); to_device( mi );
  const vpb_grid_t * hg = sp->g;
  PushCoef c;
  c.qdt_2mc = (sp->q*hg->dt)/(2*sp->m*hg->cvac);      // advance_p_pipeline.cc:267-271
  c.cdt_dx  = hg->cvac*hg->dt*hg->rdx;
  c.cdt_dy  = hg->cvac*hg->dt*hg->rdy;
  c.cdt_dz  = hg->cvac*hg->dt*hg->rdz;
  c.qsp     = sp->q;
  VPB_CUDA( cudaMemsetAsync( ms->d_counters, 0, sizeof(int)*2, r.stream ) );
  if( ms->np>0 ) {
    if( r.deposit_mode==VPB_DEPOSIT_PARITY ) advance_p_parity( ms, ma, mi, c );
    else {
      const int nblock = div_up( ms->np, AP_STRIP*AP_WARPS );
      k_advance_p<<< nblock, 32*AP_WARPS, 0, r.stream >>>( ms->d_p, ms->np, mi->d, ma->d, ms->d_pm, (int)ms->cap_pm,
                                                            ms->d_counters, ms->grid->dev, c );
      VPB_CHECK_LAUNCH(); r.launches++;
    }
  }
  // The mover count is what the host driver reads next (advance.cc:73-101).
  int cnt[2];
  VPB_CUDA( cudaMemcpyAsync( cnt, ms->d_counters, sizeof(int)*2, cudaMemcpyDeviceToHost, r.stream ) );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( cnt[1] ) VPB_WARNING( "Species \"%s\" ran out of storage for %i movers", sp->name, cnt[1] );   // advance_p_pipeline.cc:301-305
  ms->nm = cnt[0]<(int)ms->cap_pm ? cnt[0] : (int)ms->cap_pm;
  if( ms->nm>1 ) sort_movers_by_index( ms );
  mark_dev_written( ms ); mark_dev_written( ma );
  movers_to_host( ms );
  if( r.coherence==VPB_COHERENCE_STRICT ) { to_host( ms ); to_host( ma ); }
}
