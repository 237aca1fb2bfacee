// vpb_particles.cu -- energy_p, center_p / uncenter_p, the synthetic loader
// and small species helpers.
// Reference: src/species_advance/standard/pipeline/energy_p_pipeline.cc:17-115,
//            .../center_p_pipeline.cc, .../uncenter_p_pipeline.cc:16-98.
#include "vpb_push.cuh"
#include <algorithm>
#include <vector>

namespace vpb {

// ---- energy_p: 32 B read per particle, double partial sums -----------------
constexpr int EP_BLOCK = 256;

__global__ void __launch_bounds__(EP_BLOCK)
k_energy_p( double * __restrict__ partial, const vpb_particle_t * __restrict__ p, int np,
            const float * __restrict__ ip, float qdt_2mc, float msp ) {
  const float one = 1.0f;
  double en = 0.0;
  for( long long n = (long long)blockIdx.x*blockDim.x + threadIdx.x; n<np; n += (long long)gridDim.x*blockDim.x ) {
    const float4 * pp = reinterpret_cast<const float4 *>( p + n );
    const float4 r0 = __ldg( pp ), r1 = __ldg( pp+1 );
    const float dx = r0.x, dy = r0.y, dz = r0.z; const int i = __float_as_int( r0.w );
    const Interp f = load_interp( ip, i );
    float v0 = r1.x + qdt_2mc*( ( f.ex + dy*f.dexdy ) + dz*( f.dexdz + dy*f.d2exdydz ) );
    float v1 = r1.y + qdt_2mc*( ( f.ey + dz*f.deydz ) + dx*( f.deydx + dz*f.d2eydzdx ) );
    float v2 = r1.z + qdt_2mc*( ( f.ez + dx*f.dezdx ) + dy*( f.dezdy + dx*f.d2ezdxdy ) );
    v0 = ( v0*v0 + v1*v1 ) + v2*v2;
    v0 = ( msp*r1.w )*( v0/( one + sqrtf( one + v0 ) ) );
    en += (double)v0;
  }
  __shared__ double s[EP_BLOCK/32];
  for( int o=16; o; o>>=1 ) en += __shfl_down_sync( 0xffffffffu, en, o );
  if( ( threadIdx.x & 31 )==0 ) s[threadIdx.x>>5] = en;
  __syncthreads();
  if( threadIdx.x==0 ) {
    double t = 0; for( int k=0; k<EP_BLOCK/32; k++ ) t += s[k];
    partial[blockIdx.x] = t;
  }
}


// ---- center_p / uncenter_p -------------------------------------------------
template<bool UNCENTER>
__global__ void __launch_bounds__(256)
k_half_step( vpb_particle_t * __restrict__ p, int np, const float * __restrict__ ip, float args_qdt_2mc ) {
  const long long n = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  const float qdt_2mc = UNCENTER ? -args_qdt_2mc : args_qdt_2mc;
  const float qdt_4mc = UNCENTER ? (float)( -0.5*(double)args_qdt_2mc ) : (float)( 0.5*(double)args_qdt_2mc );
  const float one = 1.0f, one_third = 1.0/3.0, two_fifteenths = 2.0/15.0;
  float4 * pp = reinterpret_cast<float4 *>( p + n );
  const float4 r0 = pp[0]; float4 r1 = pp[1];
  const float dx = r0.x, dy = r0.y, dz = r0.z;
  const Interp f = load_interp( ip, __float_as_int( r0.w ) );
  const float hax = qdt_2mc*( ( f.ex + dy*f.dexdy ) + dz*( f.dexdz + dy*f.d2exdydz ) );
  const float hay = qdt_2mc*( ( f.ey + dz*f.deydz ) + dx*( f.deydx + dz*f.d2eydzdx ) );
  const float haz = qdt_2mc*( ( f.ez + dx*f.dezdx ) + dy*( f.dezdy + dx*f.d2ezdxdy ) );
  const float cbx = f.cbx + dx*f.dcbxdx, cby = f.cby + dy*f.dcbydy, cbz = f.cbz + dz*f.dcbzdz;
  float ux = r1.x, uy = r1.y, uz = r1.z, v0, v1, v2, v3, v4;
  if( !UNCENTER ) { ux += hax; uy += hay; uz += haz; }
  v0 = qdt_4mc/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
  v1 = cbx*cbx + ( cby*cby + cbz*cbz );
  v2 = ( v0*v0 )*v1;
  v3 = v0*( one + v2*( one_third + v2*two_fifteenths ) );
  v4 = v3/( one + v1*( v3*v3 ) );
  v4 += v4;
  v0 = ux + v3*( uy*cbz - uz*cby );
  v1 = uy + v3*( uz*cbx - ux*cbz );
  v2 = uz + v3*( ux*cby - uy*cbx );
  ux += v4*( v1*cbz - v2*cby );
  uy += v4*( v2*cbx - v0*cbz );
  uz += v4*( v0*cby - v1*cbx );
  if( UNCENTER ) { ux += hax; uy += hay; uz += haz; }
  r1.x = ux; r1.y = uy; r1.z = uz;
  pp[1] = r1;
}

// ---- synthetic loader (bench only; counter-based generator) ----------------
__device__ __forceinline__ uint64_t splitmix64( uint64_t x ) {
  x += 0x9E3779B97F4A7C15ull; x = ( x ^ ( x>>30 ) )*0xBF58476D1CE4E5B9ull;
  x = ( x ^ ( x>>27 ) )*0x94D049BB133111EBull; return x ^ ( x>>31 );
}
__device__ __forceinline__ float u01( uint64_t h ) { return ( (float)( h>>40 ) + 0.5f )*( 1.0f/16777216.0f ); }

__global__ void __launch_bounds__(256)
k_load_maxwellian( vpb_particle_t * __restrict__ p, long long np, int ppc, GridDev g, float vth, float w, uint64_t seed ) {
  const long long n = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  const long long cell = n/ppc;
  const int x = (int)( cell % g.nx ), y = (int)( ( cell/g.nx ) % g.ny ), z = (int)( cell/( (long long)g.nx*g.ny ) );
  const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
  const uint64_t h0 = splitmix64( seed ^ (uint64_t)n*0xD1342543DE82EF95ull );
  const uint64_t h1 = splitmix64( h0 ), h2 = splitmix64( h1 ), h3 = splitmix64( h2 ), h4 = splitmix64( h3 ), h5 = splitmix64( h4 );
  const float dx = 2.f*u01( h0 )-1.f, dy = 2.f*u01( h1 )-1.f, dz = 2.f*u01( h2 )-1.f;
  // Box-Muller
  const float r1 = sqrtf( -2.f*logf( u01( h3 ) ) ), r2 = sqrtf( -2.f*logf( u01( h4 ) ) );
  const float a1 = 6.2831853f*u01( h5 ), a2 = 6.2831853f*u01( splitmix64( h5 ) );
  float4 * pp = reinterpret_cast<float4 *>( p + n );
  pp[0] = make_float4( dx, dy, dz, __int_as_float( v ) );
  pp[1] = make_float4( vth*r1*cosf( a1 ), vth*r1*sinf( a1 ), vth*r2*cosf( a2 ), w );
}

// Langevin thermostat: u = decay u + drive N(0,1), three independent normal deviates per particle from a
// counter-based generator (splitmix64 of (key, step, particle index): any particle's deviates can be
// regenerated from its index alone, nothing to store, nothing shared between threads).
__global__ void __launch_bounds__(256)
k_langevin( vpb_particle_t * __restrict__ p, int np, float decay, float drive, uint64_t key, uint64_t step ) {
  const int n = blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  const uint64_t h0 = splitmix64( key ^ splitmix64( step*0x9e3779b97f4a7c15ull + (uint64_t)n ) );
  const uint64_t h1 = splitmix64( h0 ), h2 = splitmix64( h1 ), h3 = splitmix64( h2 );
  const float r1 = sqrtf( -2.f*logf( u01( h0 ) ) ), r2 = sqrtf( -2.f*logf( u01( h1 ) ) );
  const float a1 = 6.2831853f*u01( h2 ), a2 = 6.2831853f*u01( h3 );
  float4 * q = reinterpret_cast<float4 *>( p + n ) + 1;
  float4 u = *q;
  u.x = decay*u.x + drive*( r1*cosf( a1 ) );
  u.y = decay*u.y + drive*( r1*sinf( a1 ) );
  u.z = decay*u.z + drive*( r2*cosf( a2 ) );
  *q = u;
}

// Movers come out of advance_p in atomic-append order; the reference's order
// (single pipeline) is ascending particle index, and boundary_p's back-fill
// makes the final particle order depend on it.  Bitonic sort of the 16 B
// movers by .i on the device: chunks of MS_CHUNK in shared memory, the wide
// compare-exchange steps in global memory.
constexpr int MS_CHUNK = 2048, MS_T = 1024;

__device__ __forceinline__ void ms_cmpx( float4 & a, float4 & b, bool up ) {
  if( ( __float_as_int( a.w )>__float_as_int( b.w ) )==up ) { const float4 t = a; a = b; b = t; }
}

// mode 0: sort each chunk completely (k = 2..MS_CHUNK); mode 1: finish stage k
// inside each chunk (j = MS_CHUNK/2..1).  Direction from the GLOBAL index.
__global__ void __launch_bounds__(MS_T)
k_movers_bitonic_local( float4 * __restrict__ d, int k_stage, int mode ) {
  __shared__ float4 s[MS_CHUNK];
  const int base = blockIdx.x*MS_CHUNK;
  s[threadIdx.x] = d[base+threadIdx.x]; s[threadIdx.x+MS_T] = d[base+threadIdx.x+MS_T];
  __syncthreads();
  for( int k = mode==0 ? 2 : k_stage; k<=( mode==0 ? MS_CHUNK : k_stage ); k<<=1 ) {
    for( int j = ( k>MS_CHUNK ? MS_CHUNK : k )>>1; j>0; j>>=1 ) {
      const int t = threadIdx.x;
      const int lo = ( ( t & ~(j-1) )<<1 ) | ( t & (j-1) ), hi = lo | j;
      const bool up = ( ( base+lo ) & k )==0;
      ms_cmpx( s[lo], s[hi], up );
      __syncthreads();
    }
  }
  d[base+threadIdx.x] = s[threadIdx.x]; d[base+threadIdx.x+MS_T] = s[threadIdx.x+MS_T];
}

__global__ void __launch_bounds__(256)
k_movers_bitonic_global( float4 * __restrict__ d, int n2, int k, int j ) {
  const int t = blockIdx.x*blockDim.x + threadIdx.x;
  if( t>=n2/2 ) return;
  const int lo = ( ( t & ~(j-1) )<<1 ) | ( t & (j-1) ), hi = lo | j;
  float4 a = d[lo], b = d[hi];
  const bool up = ( lo & k )==0;
  if( ( __float_as_int( a.w )>__float_as_int( b.w ) )==up ) { d[lo] = b; d[hi] = a; }
}

__global__ void k_movers_pad( float4 * __restrict__ d, const float4 * __restrict__ src, int n, int n2 ) {
  const int t = blockIdx.x*blockDim.x + threadIdx.x;
  if( t<n2 ) d[t] = t<n ? src[t] : make_float4( 0.f, 0.f, 0.f, __int_as_float( 0x7fffffff ) );
}

void sort_movers_by_index( SpeciesMirror * m ) {
  Runtime & r = rt();
  const int n = m->nm;
  if( n<2 ) return;
  int n2 = MS_CHUNK; while( n2<n ) n2 <<= 1;
  float4 * d = (float4 *)scratch( sizeof(float4)*(size_t)n2 );
  k_movers_pad<<< div_up( n2, 256 ), 256, 0, r.stream >>>( d, (const float4 *)m->d_pm, n, n2 );
  k_movers_bitonic_local<<< n2/MS_CHUNK, MS_T, 0, r.stream >>>( d, 0, 0 );
  r.launches += 2;
  for( int k=2*MS_CHUNK; k<=n2; k<<=1 ) {
    for( int j=k>>1; j>=MS_CHUNK; j>>=1 ) { k_movers_bitonic_global<<< div_up( n2/2, 256 ), 256, 0, r.stream >>>( d, n2, k, j ); r.launches++; }
    k_movers_bitonic_local<<< n2/MS_CHUNK, MS_T, 0, r.stream >>>( d, k, 1 ); r.launches++;
  }
  VPB_CHECK_LAUNCH();
  VPB_CUDA( cudaMemcpyAsync( m->d_pm, d, sizeof(float4)*(size_t)n, cudaMemcpyDeviceToDevice, r.stream ) );
}

} // namespace vpb

using namespace vpb;

extern "C" {

double energy_p( const vpb_species_t * sp, const vpb_interpolator_array_t * ia ) {
  if( !sp || !ia || sp->g!=ia->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( const_cast<vpb_species_t *>( sp ) );
  InterpMirror  * mi = mirror( const_cast<vpb_interpolator_array_t *>( ia ) );
  to_device( ms ); to_device( mi );
  const vpb_grid_t * hg = sp->g;
  const float qdt_2mc = (sp->q*hg->dt)/(2*sp->m*hg->cvac);
  if( r.deposit_mode==VPB_DEPOSIT_PARITY ) {      // the reference's summation order, bit for bit
    double loc = energy_p_parity( ms, mi, qdt_2mc, sp->m );
    if( g_world_size>1 ) comm_allsum_d( &loc, 1 );
    return loc*( (double)hg->cvac*(double)hg->cvac );
  }
  int nb = div_up( ms->np>0 ? ms->np : 1, EP_BLOCK ); if( nb>r.sm_count*8 ) nb = r.sm_count*8;
  double * partial = (double *)scratch( sizeof(double)*( (size_t)nb + 8 ) );
  k_energy_p<<< nb, EP_BLOCK, 0, r.stream >>>( partial, ms->d_p, ms->np, mi->d, qdt_2mc, sp->m );
  k_sum_partials<<< 1, 32, 0, r.stream >>>( partial + nb, partial, nb, 1, 1.0 );
  VPB_CHECK_LAUNCH(); r.launches += 2;
  double local = 0;
  d2h( &local, partial + nb, sizeof(double), r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( g_world_size>1 ) comm_allsum_d( &local, 1 );
  return local*( (double)hg->cvac*(double)hg->cvac );
}

static void half_step( vpb_species_t * sp, const vpb_interpolator_array_t * ia, bool uncenter ) {
  if( !sp || !ia || sp->g!=ia->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( sp );
  InterpMirror  * mi = mirror( const_cast<vpb_interpolator_array_t *>( ia ) );
  to_device( ms ); to_device( mi );
  const vpb_grid_t * hg = sp->g;
  const float qdt_2mc = (sp->q*hg->dt)/(2*sp->m*hg->cvac);
  if( ms->np>0 ) {
    if( uncenter ) k_half_step<true ><<< div_up(ms->np,256), 256, 0, r.stream >>>( ms->d_p, ms->np, mi->d, qdt_2mc );
    else           k_half_step<false><<< div_up(ms->np,256), 256, 0, r.stream >>>( ms->d_p, ms->np, mi->d, qdt_2mc );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
  finish_write( ms );
}
void center_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia )   { half_step( sp, ia, false ); }
void uncenter_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia ) { half_step( sp, ia, true ); }

// ---- Langevin thermostat (src/collision/pipeline/langevin_pipeline.c:20-42, 44-112) ----------------
static uint64_t g_collision_seed = 0x5eedull;
void vpic_b200_set_collision_seed( uint64_t seed ) { g_collision_seed = seed; }

void apply_langevin_pipeline( vpb_langevin_t * l ) {
  if( !l || !l->sp || !l->sp->g ) VPB_ERROR( "Bad args" );
  vpb_species_t * sp = l->sp;
  if( l->interval<1 || ( sp->g->step % l->interval ) ) return;                   // langevin_pipeline.c:47-51
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( sp );
  to_device( ms );
  // langevin_pipeline.c:99-108 (float nudt; exp / expm1 / sqrt in double, rounded to float)
  const float nudt = l->nu*(float)l->interval*sp->g->dt;
  const float decay = (float)exp( -(double)nudt );
  const float drive = (float)sqrt( ( -expm1( -2*(double)nudt )*l->kT )/( sp->m*sp->g->cvac ) );
  if( ms->np>0 ) {
    const uint64_t key = g_collision_seed ^ ( (uint64_t)(unsigned)g_world_rank<<48 ) ^ ( (uint64_t)(unsigned)sp->id<<40 );
    k_langevin<<< div_up( ms->np, 256 ), 256, 0, r.stream >>>( ms->d_p, ms->np, decay, drive, key, (uint64_t)sp->g->step );
    VPB_CHECK_LAUNCH(); r.launches++;
  }
  const bool keys = ms->keys_valid, home = ms->home_valid;     // momenta only: voxels, order and keys are untouched
  finish_write( ms );
  ms->keys_valid = keys; ms->home_valid = home;
}

void vpic_b200_load_uniform_maxwellian( vpb_species_t * sp, int ppc, float vth, float w, uint64_t seed ) {
  if( !sp || ppc<1 ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  SpeciesMirror * ms = mirror( sp );
  const GridDev & g = ms->grid->dev;
  const long long np = (long long)ppc*g.nx*g.ny*g.nz;
  if( np>(long long)ms->cap_p ) VPB_ERROR( "species \"%s\": %lld particles do not fit max_np=%zu", sp->name, np, ms->cap_p );
  k_load_maxwellian<<< div_up(np,256), 256, 0, r.stream >>>( ms->d_p, np, ppc, g, vth, w, seed );
  VPB_CHECK_LAUNCH(); r.launches++;
  ms->np = (int)np; ms->nm = 0; sp->np = (int)np; sp->nm = 0;
  ms->partition_dev_valid = false;
  mark_dev_written( ms );
  if( r.coherence==VPB_COHERENCE_STRICT && sp->p ) to_host( ms );
}

} // extern "C"
