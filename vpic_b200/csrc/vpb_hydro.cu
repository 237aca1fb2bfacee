// vpb_hydro.cu -- hydro moments on the device: the species
// diagnostics every dump_hydro of a deck goes through.
// Reference: src/species_advance/standard/hydro_p.c (accumulate_hydro_p),
// src/sf_interface/hydro_array.c (clear / synchronize / new / delete).
//
// Device layout == host layout: 16 floats (64 B) per node, so the coherence
// copies are plain memcpys and one node's 14 moments are four 128-bit vector
// reductions (red.global.add.v4.f32).  FAST deposit: one thread per particle,
// 8 nodes x 4 vector reductions straight to L2 (the order of the float adds is
// not the reference's: tolerance).  PARITY deposit: vpb_parity.cu adds every
// node's contributions in particle order, bit-identical with the serial loop.
#include "vpb_runtime.h"
#include "vpb_hydro.cuh"

namespace vpb {

HydroMirror * mirror( vpb_hydro_array_t * ha ) {
  ensure_init();
  Runtime & r = rt();
  if( !ha ) VPB_ERROR( "NULL hydro array" );
  auto it = r.hydros.find( ha );
  if( it!=r.hydros.end() ) return it->second;
  HydroMirror * m = new HydroMirror;
  m->host = ha; m->grid = mirror( ha->g );
  const GridDev & g = m->grid->dev;
  VPB_CUDA( cudaMalloc( &m->d, sizeof(float)*HY_STRIDE*(size_t)g.nv ) );
  VPB_CUDA( cudaMemsetAsync( m->d, 0, sizeof(float)*HY_STRIDE*(size_t)g.nv, r.stream ) );
  const size_t a = (size_t)( g.ny+1 )*( g.nz+1 ), b = (size_t)( g.nz+1 )*( g.nx+1 ), c = (size_t)( g.nx+1 )*( g.ny+1 );
  m->face_floats = 14*( a>b ? ( a>c ? a : c ) : ( b>c ? b : c ) ) + 32;
  VPB_CUDA( cudaMalloc( &m->d_face, sizeof(float)*4*m->face_floats ) );
  r.hydros[ha] = m;
  return m;
}

void to_device( HydroMirror * m ) {
  Runtime & r = rt();
  if( m->dev_valid && ( r.coherence!=VPB_COHERENCE_STRICT || !m->host_valid ) ) return;
  if( !m->host->h ) { if( m->dev_valid ) return; VPB_ERROR( "hydro array has neither host nor device data" ); }
  static_assert( sizeof(vpb_hydro_t)==sizeof(float)*HY_STRIDE, "device hydro layout == host layout" );
  const size_t bytes = sizeof(vpb_hydro_t)*(size_t)m->grid->dev.nv;
  pin_host( m->host->h, bytes );
  h2d( m->d, m->host->h, bytes, r.stream );
  m->dev_valid = true;
}

void to_host( HydroMirror * m ) {
  Runtime & r = rt();
  if( !m->dev_valid || m->host_valid || !m->host->h ) return;
  d2h( m->host->h, m->d, sizeof(vpb_hydro_t)*(size_t)m->grid->dev.nv, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  m->host_valid = true;
}

void release( HydroMirror * m ) {
  unpin_host( m->host->h );
  cudaFree( m->d ); cudaFree( m->d_face );
  delete m;
}

namespace {

__global__ void __launch_bounds__(256)
k_accumulate_hydro_p( float * __restrict__ hy, const vpb_particle_t * __restrict__ p, int np,
                      const float * __restrict__ ip, int sy, int sz, HydroCoef k ) {
  const int n = blockIdx.x*blockDim.x + threadIdx.x;
  if( n>=np ) return;
  const float4 * pp = reinterpret_cast<const float4 *>( p + n );
  const float4 r0 = __ldg( pp ), r1 = __ldg( pp+1 );
  const int v = __float_as_int( r0.w );
  const HydroParticle h = hydro_particle( r0, r1, load_interp( ip, v ), k );
  const int off[8] = { 0, 1, sy, sy+1, sz, sz+1, sz+sy, sz+sy+1 };
# pragma unroll
  for( int c=0; c<8; c++ ) {
    float o[14];
    hydro_node_terms( o, h.wn[c], k, h.ux, h.uy, h.uz, h.vx, h.vy, h.vz, h.ke_mc );
    float * a = hy + (size_t)( v + off[c] )*HY_STRIDE;
    red_add_v4( a,    o[0],  o[1],  o[2],  o[3]  );
    red_add_v4( a+4,  o[4],  o[5],  o[6],  o[7]  );
    red_add_v4( a+8,  o[8],  o[9],  o[10], o[11] );
    red_add_v4( a+12, o[12], o[13], 0.f,   0.f   );
  }
}

// Node planes normal to X: u along Y (1..nY+1), w along Z (1..nZ+1); the
// message order is the reference's z,y,x nesting (hydro_array.c:60-69).
enum { HO_DOUBLE, HO_PACK, HO_FOLD };
__global__ void __launch_bounds__(128)
k_hydro_plane( float * __restrict__ hy, GridDev g, int op, int X, int plane, float lw, float rw, float * __restrict__ buf ) {
  const int n3[3] = { g.nx, g.ny, g.nz }, s3[3] = { 1, g.sy, g.sz };
  const int Y = (X+1)%3, Z = (X+2)%3;
  const int t = blockIdx.x*blockDim.x + threadIdx.x;
  // item order: ascending voxel index inside the plane == z,y,x nesting
  int a0, a1;          // the faster / slower running in-plane axis
  if( X==0 ) { a0 = 1; a1 = 2; } else if( X==1 ) { a0 = 0; a1 = 2; } else { a0 = 0; a1 = 1; }
  (void)Y; (void)Z;
  const int e0 = n3[a0]+1, e1 = n3[a1]+1;
  if( t>=e0*e1 ) return;
  const int v = plane*s3[X] + ( 1 + t%e0 )*s3[a0] + ( 1 + t/e0 )*s3[a1];
  float4 * h = reinterpret_cast<float4 *>( hy + (size_t)v*HY_STRIDE );
  float * m = buf + 14*(size_t)t;
  float4 q0 = h[0], q1 = h[1], q2 = h[2], q3 = h[3];
  float c[14] = { q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, q3.x, q3.y };
  if( op==HO_PACK ) {
#   pragma unroll
    for( int i=0; i<14; i++ ) m[i] = c[i];
    return;
  }
  if( op==HO_DOUBLE ) {
#   pragma unroll
    for( int i=0; i<14; i++ ) c[i] *= 2;
  } else {
#   pragma unroll
    for( int i=0; i<14; i++ ) c[i] = lw*c[i] + rw*m[i];
  }
  h[0] = make_float4( c[0], c[1], c[2],  c[3]  ); h[1] = make_float4( c[4],  c[5],  c[6], c[7] );
  h[2] = make_float4( c[8], c[9], c[10], c[11] ); h[3] = make_float4( c[12], c[13], q3.z, q3.w );
}

void launch_plane( HydroMirror * m, int op, int X, int plane, float lw, float rw, float * buf ) {
  Runtime & r = rt();
  const GridDev & g = m->grid->dev;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const int items = ( n3[(X+1)%3]+1 )*( n3[(X+2)%3]+1 );
  k_hydro_plane<<< div_up( items, 128 ), 128, 0, r.stream >>>( m->d, g, op, X, plane, lw, rw, buf );
  VPB_CHECK_LAUNCH(); r.launches++;
}

} // namespace
} // namespace vpb

using namespace vpb;

extern "C" {

// hydro_array.c:52-56
void clear_hydro_array( vpb_hydro_array_t * ha ) {
  if( !ha ) VPB_ERROR( "NULL hydro array" );
  Runtime & r = rt();
  HydroMirror * m = mirror( ha );
  const size_t bytes = sizeof(float)*HY_STRIDE*(size_t)m->grid->dev.nv;
  VPB_CUDA( cudaMemsetAsync( m->d, 0, bytes, r.stream ) );
  r.launches++;
  finish_write( m );
}

// hydro_p.c:26-167
void accumulate_hydro_p( vpb_hydro_array_t * ha, const vpb_species_t * sp, const vpb_interpolator_array_t * ia ) {
  if( !ha || !sp || !ia || ha->g!=sp->g || ha->g!=ia->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  HydroMirror   * m  = mirror( ha );
  SpeciesMirror * ms = mirror( const_cast<vpb_species_t *>( sp ) );
  InterpMirror  * mi = mirror( const_cast<vpb_interpolator_array_t *>( ia ) );
  to_device( m ); to_device( ms ); to_device( mi );
  const vpb_grid_t * hg = sp->g; const GridDev & g = m->grid->dev;
  HydroCoef k;
  k.c = hg->cvac; k.qsp = sp->q; k.mspc = sp->m*k.c;
  k.qdt_2mc = ( k.qsp*hg->dt )/( 2*k.mspc ); k.qdt_4mc2 = k.qdt_2mc/( 2*k.c ); k.r8V = hg->r8V;
  if( ms->np>0 ) {
    if( r.deposit_mode==VPB_DEPOSIT_PARITY ) accumulate_hydro_p_parity( m, ms, mi, k );
    else {
      k_accumulate_hydro_p<<< div_up( ms->np, 256 ), 256, 0, r.stream >>>( m->d, ms->d_p, ms->np, mi->d, g.sy, g.sz, k );
      VPB_CHECK_LAUNCH(); r.launches++;
    }
  }
  finish_write( m );
}

// hydro_array.c:71-222
void synchronize_hydro_array( vpb_hydro_array_t * ha ) {
  if( !ha ) VPB_ERROR( "NULL hydro array" );
  HydroMirror * m = mirror( ha );
  to_device( m );
  const GridMirror * gm = m->grid; const GridDev & g = gm->dev; const vpb_grid_t * hg = gm->host;
  const int n3[3] = { g.nx, g.ny, g.nz };
  const float d3[3] = { hg->dx, hg->dy, hg->dz };
  // half cells on locally handled faces (:86-121); x2 is exact, the face order does not matter
  for( int X=0; X<3; X++ ) for( int side=-1; side<=1; side+=2 )
    if( !face_is_remote( gm, X, side ) ) launch_plane( m, HO_DOUBLE, X, side<0 ? 1 : n3[X]+1, 0.f, 0.f, m->d_face );
  float * slo = m->d_face, * shi = slo + m->face_floats, * rlo = shi + m->face_floats, * rhi = rlo + m->face_floats;
  for( int X=0; X<3; X++ ) {
    const bool rem_lo = face_is_remote( gm, X, -1 ), rem_hi = face_is_remote( gm, X, +1 );
    if( !rem_lo && !rem_hi ) continue;
    const int nf = 14*( n3[(X+1)%3]+1 )*( n3[(X+2)%3]+1 );
    if( rem_lo ) launch_plane( m, HO_PACK, X, 1,       0.f, 0.f, slo );
    if( rem_hi ) launch_plane( m, HO_PACK, X, n3[X]+1, 0.f, 0.f, shi );
    const float * from_lo = nullptr, * from_hi = nullptr;
    const bool self_lo = rem_lo && face_bc( gm, X, -1 )==g_world_rank, self_hi = rem_hi && face_bc( gm, X, +1 )==g_world_rank;
    if( ( rem_lo && !self_lo ) || ( rem_hi && !self_hi ) ) {
      const size_t nb = sizeof(float)*(size_t)nf;
      comm_sendrecv( rem_lo ? gm->fbc[X] : -1, rem_hi ? gm->fbc[X+3] : -1, slo, rem_lo ? nb : 0, shi, rem_hi ? nb : 0,
                     rlo, rem_lo ? nb : 0, rhi, rem_hi ? nb : 0 );
      if( rem_lo ) from_lo = rlo;
      if( rem_hi ) from_hi = rhi;
    } else {                                   // periodic onto ourselves: our own opposite message (grid_comm.c:20-23)
      if( rem_lo ) from_lo = shi;
      if( rem_hi ) from_hi = slo;
    }
    // the neighbour's cell size travels in the message (:131); bricks of one run share it
    float rw = d3[X], lw = rw + d3[X];
    rw /= lw; lw = d3[X]/lw; lw += lw; rw += rw;
    if( from_lo ) launch_plane( m, HO_FOLD, X, 1,       lw, rw, const_cast<float *>( from_lo ) );
    if( from_hi ) launch_plane( m, HO_FOLD, X, n3[X]+1, lw, rw, const_cast<float *>( from_hi ) );
  }
  // A hydro array exists for the dump that follows (dump.cc:227-229 reads ha->h right after
  // this call): the host copy is made current in RESIDENT mode as well, so that unmodified
  // dump code works without a vpic_b200_sync_to_host.
  mark_dev_written( m ); to_host( m );
}

// Standalone construction (a reference build keeps its own new_hydro_array /
// delete_hydro_array and its checkpoint of the host struct; hydro_array.c:32-50).
vpb_hydro_array_t * vpic_b200_new_hydro_array( vpb_grid_t * g ) {
  if( !g ) VPB_ERROR( "NULL grid" );
  vpb_hydro_array_t * ha = (vpb_hydro_array_t *)calloc( 1, sizeof(*ha) );
  void * h = nullptr;
  if( !ha || posix_memalign( &h, 128, sizeof(vpb_hydro_t)*(size_t)g->nv ) ) VPB_ERROR( "out of memory" );
  memset( h, 0, sizeof(vpb_hydro_t)*(size_t)g->nv );
  ha->h = (vpb_hydro_t *)h; ha->g = g;
  return ha;
}

void vpic_b200_delete_hydro_array( vpb_hydro_array_t * ha ) {
  if( !ha ) return;
  vpic_b200_forget( ha );
  free( ha->h ); free( ha );
}

} // extern "C"
