// vpb_mirrors.cu -- interpolator / accumulator / species mirrors and the
// control-surface entry points of vpic_b200.h part 2.
#include "vpb_runtime.h"
#include <cstring>

namespace vpb {

// ---- interpolator ----------------------------------------------------------

InterpMirror * mirror( vpb_interpolator_array_t * ia ) {
  ensure_init();
  Runtime & r = rt();
  if( !ia ) VPB_ERROR( "Bad args" );
  auto it = r.interps.find( ia );
  if( it!=r.interps.end() ) return it->second;
  InterpMirror * m = new InterpMirror;
  m->host = ia; m->grid = mirror( ia->g );
  size_t bytes = sizeof(float)*IP_STRIDE*(size_t)m->grid->dev.nv;
  VPB_CUDA( cudaMalloc( &m->d, bytes ) );
  VPB_CUDA( cudaMemsetAsync( m->d, 0, bytes, r.stream ) );
  r.interps[ia] = m;
  return m;
}

// host layout (80 B records) <-> device layout (96 B records)
__global__ void k_interp_restride( float * __restrict__ dst, int dstride, const float * __restrict__ src, int sstride, int nv ) {
  const long long t = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  const long long v = t/5; const int q = (int)( t - 5*v );
  if( v>=nv ) return;
  reinterpret_cast<float4 *>( dst + v*dstride )[q] = reinterpret_cast<const float4 *>( src + v*sstride )[q];
}

void to_device( InterpMirror * m ) {
  Runtime & r = rt();
  if( m->dev_valid && ( r.coherence!=VPB_COHERENCE_STRICT || !m->host_valid ) ) return;
  if( !m->host->i ) { if( m->dev_valid ) return; VPB_ERROR( "interpolator array has neither host nor device data" ); }
  static_assert( sizeof(vpb_interpolator_t)==sizeof(float)*IP_HOST, "host interpolator layout" );
  const int nv = m->grid->dev.nv;
  if( m->dev_valid && track_clean( m->host->i, sizeof(vpb_interpolator_t)*(size_t)nv ) ) return;   // TRACKED: the host did not write it
  if( !m->d_stage ) VPB_CUDA( cudaMalloc( &m->d_stage, sizeof(float)*IP_HOST*(size_t)nv ) );
  pin_host( m->host->i, sizeof(vpb_interpolator_t)*(size_t)nv );
  h2d( m->d_stage, m->host->i, sizeof(vpb_interpolator_t)*(size_t)nv, r.stream );
  k_interp_restride<<< div_up( 5LL*nv, 256 ), 256, 0, r.stream >>>( m->d, IP_STRIDE, m->d_stage, IP_HOST, nv );
  VPB_CHECK_LAUNCH(); r.launches++;
  m->dev_valid = true;
  track_arm( m->host->i, sizeof(vpb_interpolator_t)*(size_t)nv );
}

void to_host( InterpMirror * m ) {
  Runtime & r = rt();
  if( !m->dev_valid || m->host_valid || !m->host->i ) return;
  const int nv = m->grid->dev.nv;
  if( !m->d_stage ) VPB_CUDA( cudaMalloc( &m->d_stage, sizeof(float)*IP_HOST*(size_t)nv ) );
  k_interp_restride<<< div_up( 5LL*nv, 256 ), 256, 0, r.stream >>>( m->d_stage, IP_HOST, m->d, IP_STRIDE, nv );
  VPB_CHECK_LAUNCH(); r.launches++;
  d2h( m->host->i, m->d_stage, sizeof(vpb_interpolator_t)*(size_t)nv, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  m->host_valid = true;
  track_arm( m->host->i, sizeof(vpb_interpolator_t)*(size_t)nv );
}

// ---- accumulator -----------------------------------------------------------
// The device holds ONE accumulator copy; it corresponds to the host's copy 0
// (sf_interface.h:107-113).  The per-pipeline private copies exist on the CPU
// only to avoid races; on the device the warp-reduced atomics make them
// unnecessary, and reduce_accumulator_array becomes a no-op.

AccMirror * mirror( vpb_accumulator_array_t * aa ) {
  ensure_init();
  Runtime & r = rt();
  if( !aa ) VPB_ERROR( "Bad args" );
  auto it = r.accs.find( aa );
  if( it!=r.accs.end() ) return it->second;
  AccMirror * m = new AccMirror;
  m->host = aa; m->grid = mirror( aa->g );
  size_t bytes = sizeof(float)*ACC_STRIDE*(size_t)m->grid->dev.nv;
  VPB_CUDA( cudaMalloc( &m->d, bytes ) );
  VPB_CUDA( cudaMemsetAsync( m->d, 0, bytes, r.stream ) );
  r.accs[aa] = m;
  return m;
}

void to_device( AccMirror * m ) {
  Runtime & r = rt();
  if( m->dev_valid && ( r.coherence!=VPB_COHERENCE_STRICT || !m->host_valid ) ) return;
  if( !m->host->a ) { if( m->dev_valid ) return; VPB_ERROR( "accumulator array has neither host nor device data" ); }
  static_assert( sizeof(vpb_accumulator_t)==sizeof(float)*ACC_STRIDE, "device accumulator layout == host layout" );
  const size_t bytes = sizeof(vpb_accumulator_t)*(size_t)m->grid->dev.nv;
  if( m->dev_valid && track_clean( m->host->a, bytes ) ) return;
  pin_host( m->host->a, bytes );
  h2d( m->d, m->host->a, bytes, r.stream );
  m->dev_valid = true;
  track_arm( m->host->a, bytes );
}

void to_host( AccMirror * m ) {
  Runtime & r = rt();
  if( !m->dev_valid || m->host_valid || !m->host->a ) return;
  d2h( m->host->a, m->d, sizeof(vpb_accumulator_t)*(size_t)m->grid->dev.nv, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  m->host_valid = true;
  track_arm( m->host->a, sizeof(vpb_accumulator_t)*(size_t)m->grid->dev.nv );
}

// ---- species ---------------------------------------------------------------

SpeciesMirror * mirror( vpb_species_t * sp ) {
  ensure_init();
  Runtime & r = rt();
  if( !sp ) VPB_ERROR( "Bad args" );
  SpeciesMirror * m;
  auto it = r.species.find( sp );
  if( it!=r.species.end() ) m = it->second;
  else {
    m = new SpeciesMirror;
    m->host = sp; m->grid = mirror( sp->g );
    VPB_CUDA( cudaMalloc( &m->d_partition, sizeof(int)*( (size_t)m->grid->dev.nv + 1 ) ) );
    VPB_CUDA( cudaMalloc( &m->d_counters, sizeof(int)*64 ) );
    VPB_CUDA( cudaMemsetAsync( m->d_counters, 0, sizeof(int)*64, r.stream ) );
    m->np = sp->np; m->nm = sp->nm;
    r.species[sp] = m;
  }
  // The host may have re-allocated p / pm (boundary_p.cc:346-394): follow max_np.
  if( (size_t)sp->max_np>m->cap_p ) {
    vpb_particle_t * n; size_t cap = (size_t)sp->max_np;
    VPB_CUDA( cudaMalloc( &n, sizeof(vpb_particle_t)*cap ) );
    if( m->d_p ) {
      materialize_sort( m );
      if( m->dev_valid && m->np>0 )
        VPB_CUDA( cudaMemcpyAsync( n, m->d_p, sizeof(vpb_particle_t)*(size_t)m->np, cudaMemcpyDeviceToDevice, r.stream ) );
      VPB_CUDA( cudaStreamSynchronize( r.stream ) );
      VPB_CUDA( cudaFree( m->d_p ) );
    }
    m->d_p = n; m->cap_p = cap;
  }
  if( (size_t)sp->max_nm>m->cap_pm ) {
    if( m->d_pm ) { VPB_CUDA( cudaStreamSynchronize( r.stream ) ); VPB_CUDA( cudaFree( m->d_pm ) ); }
    m->cap_pm = (size_t)sp->max_nm;
    VPB_CUDA( cudaMalloc( &m->d_pm, sizeof(vpb_particle_mover_t)*m->cap_pm ) );
  }
  return m;
}

void to_device( SpeciesMirror * m, bool keep_lazy_sort ) {
  Runtime & r = rt();
  if( m->dev_valid && ( r.coherence!=VPB_COHERENCE_STRICT || !m->host_valid ) ) {
    if( !keep_lazy_sort ) materialize_sort( m );
    return;
  }
  if( !m->host->p ) { if( m->dev_valid ) return; VPB_ERROR( "species \"%s\" has neither host nor device particles", m->host->name ); }
  if( m->dev_valid && m->np==m->host->np && m->nm==m->host->nm && m->nm==0 &&
      track_clean( m->host->p, sizeof(vpb_particle_t)*(size_t)m->host->np ) ) {    // TRACKED: the host did not write the particles
    if( !keep_lazy_sort ) materialize_sort( m );
    return;
  }
  m->perm_pending = false; m->keys_valid = m->keys_patchable = false; m->home_valid = false;      // the host copy replaces whatever the device held
  m->np = m->host->np; m->nm = m->host->nm;
  pin_host( m->host->p, sizeof(vpb_particle_t)*(size_t)m->host->max_np );
  if( m->np>0 )
    h2d( m->d_p, m->host->p, sizeof(vpb_particle_t)*(size_t)m->np, r.stream );
  // pending movers travel with the particles (the host may have added some:
  // inject_particle, emitters; src/vpic/misc.cc:86-99)
  if( m->nm>(int)m->cap_pm ) VPB_ERROR( "species \"%s\": nm=%d exceeds max_nm=%zu", m->host->name, m->nm, m->cap_pm );
  if( m->nm>0 && m->host->pm ) {
    h2d( m->d_pm, m->host->pm, sizeof(vpb_particle_mover_t)*(size_t)m->nm, r.stream );
    m->movers_sorted = false;
    for( int k=1; k<m->nm; k++ ) if( m->host->pm[k].i<=m->host->pm[k-1].i ) goto unsorted;
    m->movers_sorted = true;
    unsorted:;
  }
  m->dev_valid = true; m->partition_dev_valid = false;
  track_arm( m->host->p, sizeof(vpb_particle_t)*(size_t)m->host->max_np );
}

void to_host( SpeciesMirror * m ) {
  Runtime & r = rt();
  if( !m->dev_valid || m->host_valid ) return;
  materialize_sort( m );
  m->host->np = m->np;
  if( m->host->p && m->np>0 )
    d2h( m->host->p, m->d_p, sizeof(vpb_particle_t)*(size_t)m->np, r.stream );
  if( m->host->partition && m->partition_dev_valid )
    d2h( m->host->partition, m->d_partition, sizeof(int)*( (size_t)m->grid->dev.nv + 1 ), r.stream );   // nv+1 entries (species_advance.c:138); [nv] = np
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( m->host->p ) { m->host_valid = true; track_arm( m->host->p, sizeof(vpb_particle_t)*(size_t)m->host->max_np ); }
}

// The mover list and its count are what the host driver reads after advance_p
// (src/vpic/advance.cc:78-101), so they travel in every coherence mode.
void movers_to_host( SpeciesMirror * m ) {
  Runtime & r = rt();
  m->host->nm = m->nm;
  if( m->host->pm && m->nm>0 )
    d2h( m->host->pm, m->d_pm, sizeof(vpb_particle_mover_t)*(size_t)m->nm, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
}

} // namespace vpb

using namespace vpb;

// ---- control surface (vpic_b200.h part 2) ----------------------------------

extern "C" {

// world_rank / world_size as the reference's src/util/mp reports them; pure host
// state (the grid partitioners use it), no CUDA needed.
void vpic_b200_set_world( int world_rank, int world_size ) {
  g_world_size = world_size<1 ? 1 : world_size;
  g_world_rank = world_rank<0 ? 0 : world_rank;
  g_world_explicit = true;
}

int vpic_b200_init( int device, int world_rank, int world_size ) {
  Runtime & r = rt();
  if( !r.initialised ) r.device = device<0 ? 0 : device;
  g_world_rank = world_rank; g_world_size = world_size<1 ? 1 : world_size;
  g_world_explicit = true;
  ensure_init();
  return r.sm_count;
}

void vpic_b200_set_coherence( int mode ) {
  if( mode!=VPB_COHERENCE_STRICT && mode!=VPB_COHERENCE_RESIDENT && mode!=VPB_COHERENCE_TRACKED ) VPB_ERROR( "Bad args" );
  // TRACKED is STRICT for every entry point; the difference lives in to_device (skips clean arrays)
  rt().coherence = mode==VPB_COHERENCE_TRACKED ? VPB_COHERENCE_STRICT : mode;
  track_enable( mode==VPB_COHERENCE_TRACKED );
}

void vpic_b200_set_deposit_mode( int mode ) {
  if( mode!=VPB_DEPOSIT_FAST && mode!=VPB_DEPOSIT_PARITY ) VPB_ERROR( "Bad args" );
  rt().deposit_mode = mode;
}

int vpic_b200_abi_check( int s_interp, int s_acc, int s_field, int s_particle, int s_grid, int s_species ) {
  int bad = 0;
  if( s_interp  !=(int)sizeof(vpb_interpolator_t) ) bad |= 1;   // USE_V8/V16 padded layout
  if( s_acc     !=(int)sizeof(vpb_accumulator_t)  ) bad |= 2;
  if( s_field   !=(int)sizeof(vpb_field_t)        ) bad |= 4;
  if( s_particle!=(int)sizeof(vpb_particle_t)     ) bad |= 8;
  if( s_grid    !=(int)sizeof(vpb_grid_t)         ) bad |= 16;
  if( s_species !=(int)sizeof(vpb_species_t)      ) bad |= 32;
  return bad;
}

void vpic_b200_device_synchronize( void ) {
  ensure_init();
  VPB_CUDA( cudaStreamSynchronize( rt().stream ) );
}

long long vpic_b200_launch_count( void ) { return rt().launches; }
void vpic_b200_transfer_bytes( long long * h2d_d2h ) { if( h2d_d2h ) { h2d_d2h[0] = rt().h2d_bytes; h2d_d2h[1] = rt().d2h_bytes; } }
void * vpic_b200_stream( void ) { ensure_init(); return (void *)rt().stream; }

#define VPB_FIND( map, obj ) ( rt().map.find( obj )!=rt().map.end() ? rt().map[obj] : nullptr )

void vpic_b200_sync_to_host( void * obj ) {
  ensure_init();
  if( auto m = VPB_FIND( species, obj ) ) { to_host( m ); movers_to_host( m ); return; }
  if( auto m = VPB_FIND( fields,  obj ) ) { to_host( m ); return; }
  if( auto m = VPB_FIND( interps, obj ) ) { to_host( m ); return; }
  if( auto m = VPB_FIND( accs,    obj ) ) { to_host( m ); return; }
  if( auto m = VPB_FIND( hydros,  obj ) ) { to_host( m ); return; }
}

void vpic_b200_sync_from_host( void * obj ) {
  ensure_init();
  Runtime & r = rt();
  if( r.species.count( obj ) ) { auto m = mirror( (vpb_species_t *)obj ); m->dev_valid = false; m->host_valid = true; to_device( m ); return; }
  if( r.fields.count( obj )  ) { auto m = r.fields[obj];  m->dev_valid = false; m->host_valid = true; to_device( m ); return; }
  if( r.interps.count( obj ) ) { auto m = r.interps[obj]; m->dev_valid = false; m->host_valid = true; to_device( m ); return; }
  if( r.accs.count( obj )    ) { auto m = r.accs[obj];    m->dev_valid = false; m->host_valid = true; to_device( m ); return; }
  if( r.hydros.count( obj )  ) { auto m = r.hydros[obj];  m->dev_valid = false; m->host_valid = true; to_device( m ); return; }
  VPB_ERROR( "vpic_b200_sync_from_host: object %p has no device mirror yet", obj );
}

void vpic_b200_forget( void * obj ) {
  Runtime & r = rt();
  if( !r.initialised ) return;
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  if( r.species.count( obj ) ) { auto m = r.species[obj]; unpin_host( m->host->p ); release_lazy_sort( m ); cudaFree( m->d_p ); cudaFree( m->d_pm ); cudaFree( m->d_partition ); cudaFree( m->d_partition_alt ); cudaFree( m->d_counters ); delete m; r.species.erase( obj ); return; }
  if( r.fields.count( obj )  ) { auto m = r.fields[obj]; unpin_host( m->host->f ); cudaFree( m->d ); cudaFree( m->d_mat ); cudaFree( m->d_mc ); cudaFree( m->d_stage ); cudaFree( m->d_en ); cudaFree( m->d_face ); delete m; r.fields.erase( obj ); return; }
  if( r.interps.count( obj ) ) { auto m = r.interps[obj]; unpin_host( m->host->i ); cudaFree( m->d ); cudaFree( m->d_stage ); delete m; r.interps.erase( obj ); return; }
  if( r.accs.count( obj )    ) { auto m = r.accs[obj]; unpin_host( m->host->a ); cudaFree( m->d ); cudaFree( m->d_priv ); delete m; r.accs.erase( obj ); return; }
  if( r.hydros.count( obj )  ) { release( r.hydros[obj] ); r.hydros.erase( obj ); return; }
  if( r.grids.count( obj )   ) { auto m = r.grids[obj]; cudaFree( m->d_neighbor ); delete m; r.grids.erase( obj ); return; }
}

void vpic_b200_grid_changed( const void * grid ) {
  Runtime & r = rt();
  if( !r.initialised || !grid ) return;
  auto it = r.grids.find( grid );
  if( it==r.grids.end() ) return;
  GridMirror * m = it->second; const vpb_grid_t * g = m->host;
  if( m->dev.nx!=g->nx || m->dev.ny!=g->ny || m->dev.nz!=g->nz )
    VPB_ERROR( "the grid was resized (%dx%dx%d -> %dx%dx%d) while arrays on it exist", m->dev.nx, m->dev.ny, m->dev.nz, g->nx, g->ny, g->nz );
  build_neighbors( m ); m->neighbors_final = true;
}

void * vpic_b200_device_ptr( void * obj, int which ) {
  if( auto m = VPB_FIND( species, obj ) ) materialize_sort( m );
  if( auto m = VPB_FIND( species, obj ) ) return which==0 ? (void *)m->d_p : which==1 ? (void *)m->d_pm : (void *)m->d_partition;
  if( auto m = VPB_FIND( fields,  obj ) ) return (void *)m->d;
  if( auto m = VPB_FIND( interps, obj ) ) return (void *)m->d;
  if( auto m = VPB_FIND( accs,    obj ) ) return (void *)m->d;
  if( auto m = VPB_FIND( hydros,  obj ) ) return (void *)m->d;
  return nullptr;
}

// Give back the grow-on-demand device caches (sort_p's spare particle buffer, the
// scratch arena, boundary_p's message buffers, the PARITY record buffers).  Mirrors
// stay; the caches are re-grown by the next call that needs them.
void vpic_b200_trim( void ) {
  Runtime & r = rt();
  if( !r.initialised ) return;
  cudaStreamSynchronize( r.stream );
  release_boundary_buffers(); release_sort_buffers(); release_parity_buffers();
  if( r.d_scratch ) cudaFree( r.d_scratch ); r.d_scratch = nullptr; r.scratch_bytes = 0;
}

void vpic_b200_finalize( void ) {
  Runtime & r = rt();
  if( !r.initialised ) return;
  cudaStreamSynchronize( r.stream );
  std::vector<const void *> keys;
  for( auto & kv : r.species ) keys.push_back( kv.first );
  for( auto & kv : r.fields  ) keys.push_back( kv.first );
  for( auto & kv : r.interps ) keys.push_back( kv.first );
  for( auto & kv : r.accs    ) keys.push_back( kv.first );
  for( auto & kv : r.hydros  ) keys.push_back( kv.first );
  for( auto & kv : r.grids   ) keys.push_back( kv.first );
  for( auto k : keys ) vpic_b200_forget( (void *)k );
  release_boundary_buffers(); release_sort_buffers(); release_parity_buffers(); comm_shutdown(); unpin_all();
  if( r.d_scratch ) cudaFree( r.d_scratch ); r.d_scratch = nullptr; r.scratch_bytes = 0;
  if( r.h_pinned ) cudaFreeHost( r.h_pinned ); r.h_pinned = nullptr; r.pinned_bytes = 0;
}

} // extern "C"
