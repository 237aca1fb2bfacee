// vpb_runtime.cu -- runtime singleton, mirror registry, host<->device coherence.
#include "vpb_runtime.h"
#include <cstdarg>
#include <cstring>

// A host program built from the reference carries its rank / size and its
// message-passing layer (src/util/util_base.h:291-295, src/util/mp/mp.h:76-79);
// weak, so that the library also loads on its own.  An unmodified deck never
// calls vpic_b200_init / vpic_b200_comm_init: the library then takes rank and
// size from these and bootstraps its NCCL communicator over the host program's
// own allgather (vpb_comm.cu: need_comm).
extern "C" {
extern int _world_rank __attribute__((weak));
extern int _world_size __attribute__((weak));
}

namespace vpb {

int g_world_rank = 0, g_world_size = 1;
bool g_world_explicit = false;     // vpic_b200_init / vpic_b200_set_world were called

void fatal( const char * file, int line, const char * fmt, ... ) {
  // Reference convention: src/util/util_base.h:267-273
  fprintf( stderr, "Error at %s(%d)[%d]:\n\t", file, line, g_world_rank );
  va_list ap; va_start( ap, fmt ); vfprintf( stderr, fmt, ap ); va_end( ap );
  fprintf( stderr, "\n" ); fflush( stderr );
  exit( 1 );
}
void warning( const char * file, int line, const char * fmt, ... ) {
  fprintf( stderr, "Warning at %s(%d)[%d]:\n\t", file, line, g_world_rank );
  va_list ap; va_start( ap, fmt ); vfprintf( stderr, fmt, ap ); va_end( ap );
  fprintf( stderr, "\n" ); fflush( stderr );
}

Runtime & rt() { static Runtime r; return r; }

void ensure_init() {
  Runtime & r = rt();
  if( r.initialised ) return;
  int n = 0;
  bool from_host_program = false;
  if( !g_world_explicit && &_world_size && &_world_rank && _world_size>1 ) {
    g_world_rank = _world_rank; g_world_size = _world_size; from_host_program = true;
  }
  cudaError_t e = cudaGetDeviceCount( &n );
  if( e!=cudaSuccess || n<1 )
    VPB_ERROR( "libvpic_b200 needs a CUDA device (B200, sm_100a); none is usable (%s). "
               "There is no CPU fallback.", e==cudaSuccess ? "device count is 0" : cudaGetErrorString(e) );
  if( from_host_program ) {          // one process per GPU: LOCAL_RANK if the launcher gives one, else rank modulo devices
    const char * lr = getenv( "LOCAL_RANK" );
    r.device = lr ? atoi( lr ) : g_world_rank;
  }
  if( r.device>=n ) r.device = r.device % n;
  VPB_CUDA( cudaSetDevice( r.device ) );
  cudaDeviceProp prop; VPB_CUDA( cudaGetDeviceProperties( &prop, r.device ) );
  r.sm_count = prop.multiProcessorCount;
  VPB_CUDA( cudaStreamCreateWithFlags( &r.stream, cudaStreamNonBlocking ) );
  VPB_CUDA( cudaStreamCreateWithFlags( &r.copy_stream[0], cudaStreamNonBlocking ) );
  VPB_CUDA( cudaStreamCreateWithFlags( &r.copy_stream[1], cudaStreamNonBlocking ) );
  // A host program that links the library in place of the reference's own
  // kernels never calls the control surface: the same switches as environment
  // variables.
  if( const char * e = getenv( "VPIC_B200_COHERENCE" ) ) {
    if( !strcmp( e, "resident" ) ) r.coherence = VPB_COHERENCE_RESIDENT;
    else if( !strcmp( e, "strict" ) ) r.coherence = VPB_COHERENCE_STRICT;
    else if( !strcmp( e, "tracked" ) ) { r.coherence = VPB_COHERENCE_STRICT; track_enable( true ); }
    else VPB_ERROR( "VPIC_B200_COHERENCE must be strict, tracked or resident, not \"%s\"", e );
  }
  if( const char * e = getenv( "VPIC_B200_DEPOSIT" ) ) {
    if( !strcmp( e, "parity" ) ) r.deposit_mode = VPB_DEPOSIT_PARITY;
    else if( !strcmp( e, "fast" ) ) r.deposit_mode = VPB_DEPOSIT_FAST;
    else VPB_ERROR( "VPIC_B200_DEPOSIT must be fast or parity, not \"%s\"", e );
  }
  if( getenv( "VPIC_B200_REPORT" ) )      // one line at exit: proof of where the work ran
    atexit( [](){ Runtime & q = rt();
                  fprintf( stderr, "[vpic_b200] device %d: %lld kernel launches, %lld bytes host->device, %lld bytes device->host, coherence %s, deposit %s\n",
                           q.device, q.launches, q.h2d_bytes, q.d2h_bytes,
                           q.coherence==VPB_COHERENCE_STRICT ? "strict" : "resident",
                           q.deposit_mode==VPB_DEPOSIT_PARITY ? "parity" : "fast" ); } );
  r.initialised = true;
}

// The reference allocates its arrays with posix_memalign (pageable).  Page-
// locking them once lets every later cudaMemcpyAsync run at full PCIe rate and
// overlap with kernels.  Arrays below 1 MB are not worth a registration.
void pin_host( const void * ptr, size_t bytes ) {
  Runtime & r = rt();
  if( !ptr || bytes<( 1u<<20 ) ) return;
  auto it = r.pinned_host.find( ptr );
  if( it!=r.pinned_host.end() ) {
    if( it->second>=bytes ) return;
    cudaHostUnregister( const_cast<void *>( ptr ) ); r.pinned_host.erase( it );
  }
  // Page-locking costs about a second per GB (0.8 s for the 1.7 GB particle array of the LPI deck, against 0.1 s for
  // the pageable copy it was meant to speed up).  In RESIDENT coherence an array typically crosses the bus once (the
  // deck's initial state) or now and then (a dump): the first transfer of an array goes pageable, a second one
  // says the host program syncs this array repeatedly and registers it.  STRICT / TRACKED move the arrays every step.
  if( r.coherence==VPB_COHERENCE_RESIDENT && r.host_touches[ptr]++<1 ) return;
  track_disarm( ptr );               // registration wants writable pages
  if( cudaHostRegister( const_cast<void *>( ptr ), bytes, cudaHostRegisterDefault )==cudaSuccess ) r.pinned_host[ptr] = bytes;
  else (void)cudaGetLastError();     // already registered by the host program, or not lockable: copies still work
}
void unpin_host( const void * ptr ) {
  Runtime & r = rt();
  track_disarm( ptr );
  r.host_touches.erase( ptr );
  auto it = r.pinned_host.find( ptr );
  if( it==r.pinned_host.end() ) return;
  if( cudaHostUnregister( const_cast<void *>( ptr ) )!=cudaSuccess ) (void)cudaGetLastError();
  r.pinned_host.erase( it );
}
void unpin_all() {
  Runtime & r = rt();
  for( auto & kv : r.pinned_host ) if( cudaHostUnregister( const_cast<void *>( kv.first ) )!=cudaSuccess ) (void)cudaGetLastError();
  r.pinned_host.clear();
}

void * scratch( size_t bytes ) {
  Runtime & r = rt();
  if( bytes>r.scratch_bytes ) {
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    if( r.d_scratch ) VPB_CUDA( cudaFree( r.d_scratch ) );
    size_t want = bytes + bytes/16 + 4096;
    VPB_CUDA( cudaMalloc( &r.d_scratch, want ) );
    r.scratch_bytes = want;
  }
  return r.d_scratch;
}

void * pinned( size_t bytes ) {
  Runtime & r = rt();
  if( bytes>r.pinned_bytes ) {
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    if( r.h_pinned ) VPB_CUDA( cudaFreeHost( r.h_pinned ) );
    size_t want = bytes + bytes/16 + 4096;
    VPB_CUDA( cudaMallocHost( &r.h_pinned, want ) );
    r.pinned_bytes = want;
  }
  return r.h_pinned;
}

// ---- grid mirror -----------------------------------------------------------

void build_neighbors( GridMirror * m ) {
  const vpb_grid_t * g = m->host;
  GridDev & d = m->dev;
  if( m->d_neighbor ) { VPB_CUDA( cudaStreamSynchronize( rt().stream ) ); VPB_CUDA( cudaFree( m->d_neighbor ) ); m->d_neighbor = nullptr; }
  d.neighbor = nullptr;
  m->per_voxel = m->any_absorb = m->any_custom = false;
  static const int bnd[6] = { VPB_BOUNDARY(-1,0,0), VPB_BOUNDARY(0,-1,0), VPB_BOUNDARY(0,0,-1),
                              VPB_BOUNDARY( 1,0,0), VPB_BOUNDARY(0, 1,0), VPB_BOUNDARY(0,0, 1) };
  for( int f=0; f<6; f++ ) m->fbc[f] = g->bc[ bnd[f] ];

  // Classify the particle boundary of each face from g->neighbor
  // (src/grid/ops.c:58-86, 135-210) and check that the table is the regular
  // box the arithmetic neighbour computation assumes.
  const int n[3] = { g->nx, g->ny, g->nz };
  const int64_t s[3] = { 1, d.sy, d.sz };
  bool regular = g->neighbor!=nullptr;
  if( !g->neighbor ) VPB_ERROR( "grid_t.neighbor is NULL" );
  if( !g->range ) VPB_ERROR( "grid_t.range is NULL" );
  const int64_t rangem = g->range[g_world_size];
  for( int f=0; f<6; f++ ) { m->pkind[f] = PK_LOCAL; m->nb_n[f] = 0; m->pcode[f] = 0; }
  for( int f=0; f<6 && regular; f++ ) {
    const int a = f%3, side = f<3 ? -1 : 1;
    int cls = -1, kind = PK_LOCAL;
    for( int z=1; z<=n[2] && regular; z++ ) for( int y=1; y<=n[1] && regular; y++ ) for( int x=1; x<=n[0]; x++ ) {
      const int c[3] = { x, y, z };
      const int64_t v = x + s[1]*y + s[2]*z;
      const int64_t nb = g->neighbor[ 6*v + f ];
      const bool on_face = side<0 ? c[a]==1 : c[a]==n[a];
      if( !on_face ) { if( nb != g->rangel + v + side*s[a] ) { regular = false; break; } continue; }
      int k, pk = PK_LOCAL;
      if( nb==VPB_REFLECT_PARTICLES ) k = FACE_REFLECT;
      else if( nb>=g->rangel && nb<=g->rangeh ) {
        k = FACE_WRAP;
        if( nb != g->rangel + v - side*s[a]*(n[a]-1) ) { regular = false; break; }
      } else {
        k = FACE_EXIT;
        pk = nb==VPB_ABSORB_PARTICLES ? PK_ABSORB : ( nb>=0 && nb<=rangem ) ? PK_REMOTE : PK_CUSTOM;
      }
      if( cls<0 ) { cls = k; kind = pk; m->pcode[f] = pk==PK_CUSTOM ? (long long)nb : 0; }
      else if( cls!=k || kind!=pk || ( pk==PK_CUSTOM && m->pcode[f]!=(long long)nb ) ) { regular = false; break; }
    }
    d.face[f] = cls<0 ? FACE_REFLECT : cls;
    m->pkind[f] = kind;
    if( kind==PK_REMOTE ) {
      const int rank = g->bc[ bnd[f] ];
      if( rank<0 || rank>=g_world_size ) { regular = false; break; }
      const int64_t rnc = g->range[rank+1] - g->range[rank];
      const int64_t plane = (int64_t)( n[(a+1)%3]+2 )*( n[(a+2)%3]+2 );
      if( rnc%plane ) VPB_ERROR( "the brick of rank %d does not match this rank's face %d", rank, f );
      m->nb_n[f] = (int)( rnc/plane ) - 2;
    }
  }
  if( !regular ) {
    // Boundary codes inside the box (deck/wrapper.h:255-310 set_region_bc,
    // grid/ops.c set_pbc on part of a face ...): keep the table itself, 32-bit.
    // A particle can then leave from any voxel; what happens to it is decided
    // per particle from the code (boundary_p.cc:196-260).
    std::vector<int32_t> t( 6*(size_t)d.nv, NB_UNKNOWN );
    for( int f=0; f<6; f++ ) { d.face[f] = FACE_EXIT; m->pkind[f] = PK_CUSTOM; m->pcode[f] = 0; m->nb_n[f] = 0; }
    for( int z=1; z<=n[2]; z++ ) for( int y=1; y<=n[1]; y++ ) for( int x=1; x<=n[0]; x++ ) {
      const int c[3] = { x, y, z };
      const int64_t v = x + s[1]*y + s[2]*z;
      for( int f=0; f<6; f++ ) {
        const int a = f%3, side = f<3 ? -1 : 1;
        const int64_t nb = g->neighbor[ 6*v + f ];
        int32_t code;
        if( nb==VPB_REFLECT_PARTICLES ) code = NB_REFLECT;
        else if( nb==VPB_ABSORB_PARTICLES ) { code = NB_ABSORB; m->any_absorb = true; }
        else if( nb>=g->rangel && nb<=g->rangeh ) code = (int32_t)( nb - g->rangel );
        else if( nb<0 ) { code = nb>(int64_t)NB_UNKNOWN ? (int32_t)nb : NB_UNKNOWN; m->any_custom = true; }
        else if( nb<=rangem && ( side<0 ? c[a]==1 : c[a]==n[a] ) ) {
          // another rank's voxel: only through the face that rank sits behind
          const int rank = g->bc[ bnd[f] ];
          if( rank<0 || rank>=g_world_size || nb<g->range[rank] || nb>=g->range[rank+1] ) {
            // The voxel behind this face belongs to a rank the face's field boundary does not name: a deck
            // that made its box periodic, then turned the end faces into field walls, and paints the particle
            // walls later (sample/lpi_2d_F6_test:448-462, then set_region_bc) passes through this state.  The
            // reference's boundary_p could not route such a particle either (boundary_p.cc:215-230 indexes
            // range[] with bc[face]); it is dropped with the "unknown boundary interaction" warning if one
            // ever arrives here.  The table is read again at the first particle call.
            code = NB_UNKNOWN;
          } else {
            const int64_t rnc = g->range[rank+1] - g->range[rank];
            const int64_t plane = (int64_t)( n[(a+1)%3]+2 )*( n[(a+2)%3]+2 );
            if( rnc%plane ) VPB_ERROR( "the brick of rank %d does not match this rank's face %d", rank, f );
            m->nb_n[f] = (int)( rnc/plane ) - 2; m->pkind[f] = PK_REMOTE;
            code = NB_REMOTE;
          }
        } else code = NB_UNKNOWN;
        t[ 6*(size_t)v + f ] = code;
      }
    }
    VPB_CUDA( cudaMalloc( &m->d_neighbor, t.size()*sizeof(int32_t) ) );
    VPB_CUDA( cudaMemcpy( m->d_neighbor, t.data(), t.size()*sizeof(int32_t), cudaMemcpyHostToDevice ) );
    d.neighbor = m->d_neighbor;
    m->per_voxel = true;
  }
}

GridMirror * mirror( const vpb_grid_t * g ) {
  ensure_init();
  Runtime & r = rt();
  if( !g ) VPB_ERROR( "NULL grid" );
  auto it = r.grids.find( g );
  if( it!=r.grids.end() ) {
    GridMirror * m = it->second;
    if( m->dev.nx==g->nx && m->dev.ny==g->ny && m->dev.nz==g->nz ) return m;
    if( m->d_neighbor ) cudaFree( m->d_neighbor );
    delete m; r.grids.erase( it );
  }
  GridMirror * m = new GridMirror;
  m->host = g;
  GridDev & d = m->dev;
  d.nx = g->nx; d.ny = g->ny; d.nz = g->nz;
  d.sx = 1; d.sy = g->nx+2; d.sz = (g->nx+2)*(g->ny+2); d.nv = d.sz*(g->nz+2);
  fast_div_magic( (unsigned)d.sy, d.m_sy, d.sh_sy ); fast_div_magic( (unsigned)( g->ny+2 ), d.m_ny2, d.sh_ny2 );
  fast_div_magic( (unsigned)d.sz, d.m_sz, d.sh_sz );
  d.nvp = ( d.nv + 31 ) & ~31;
  d.neighbor = nullptr;
  if( d.nv!=g->nv ) VPB_ERROR( "grid_t.nv (%d) does not match (nx+2)(ny+2)(nz+2) (%d)", g->nv, d.nv );
  build_neighbors( m );
  r.grids[g] = m;
  return m;
}


// ---- field mirror ----------------------------------------------------------

__global__ void k_field_aos_to_soa( const vpb_field_t * __restrict__ aos, float * __restrict__ soa,
                                    int16_t * __restrict__ mat, int nv, int nvp ) {
  int v = blockIdx.x*blockDim.x + threadIdx.x;
  if( v>=nv ) return;
  const float4 * src = reinterpret_cast<const float4 *>( aos + v );
  float4 q0 = src[0], q1 = src[1], q2 = src[2], q3 = src[3];
  soa[ 0*(size_t)nvp+v] = q0.x; soa[ 1*(size_t)nvp+v] = q0.y; soa[ 2*(size_t)nvp+v] = q0.z; soa[ 3*(size_t)nvp+v] = q0.w;
  soa[ 4*(size_t)nvp+v] = q1.x; soa[ 5*(size_t)nvp+v] = q1.y; soa[ 6*(size_t)nvp+v] = q1.z; soa[ 7*(size_t)nvp+v] = q1.w;
  soa[ 8*(size_t)nvp+v] = q2.x; soa[ 9*(size_t)nvp+v] = q2.y; soa[10*(size_t)nvp+v] = q2.z; soa[11*(size_t)nvp+v] = q2.w;
  soa[12*(size_t)nvp+v] = q3.x; soa[13*(size_t)nvp+v] = q3.y; soa[14*(size_t)nvp+v] = q3.z; soa[15*(size_t)nvp+v] = q3.w;
  if( mat ) {
    const int16_t * ms = reinterpret_cast<const int16_t *>( aos + v ) + 32;
#   pragma unroll
    for( int c=0; c<8; c++ ) mat[ c*(size_t)nvp+v ] = ms[c];
  }
}

__global__ void k_field_soa_to_aos( vpb_field_t * __restrict__ aos, const float * __restrict__ soa, int nv, int nvp ) {
  int v = blockIdx.x*blockDim.x + threadIdx.x;
  if( v>=nv ) return;
  float4 * dst = reinterpret_cast<float4 *>( aos + v );   // material ids (last 16 B) untouched
  dst[0] = make_float4( soa[ 0*(size_t)nvp+v], soa[ 1*(size_t)nvp+v], soa[ 2*(size_t)nvp+v], soa[ 3*(size_t)nvp+v] );
  dst[1] = make_float4( soa[ 4*(size_t)nvp+v], soa[ 5*(size_t)nvp+v], soa[ 6*(size_t)nvp+v], soa[ 7*(size_t)nvp+v] );
  dst[2] = make_float4( soa[ 8*(size_t)nvp+v], soa[ 9*(size_t)nvp+v], soa[10*(size_t)nvp+v], soa[11*(size_t)nvp+v] );
  dst[3] = make_float4( soa[12*(size_t)nvp+v], soa[13*(size_t)nvp+v], soa[14*(size_t)nvp+v], soa[15*(size_t)nvp+v] );
}

FieldMirror * mirror( vpb_field_array_t * fa ) {
  ensure_init();
  Runtime & r = rt();
  if( !fa ) VPB_ERROR( "Bad args" );
  auto it = r.fields.find( fa );
  if( it!=r.fields.end() ) return it->second;
  FieldMirror * m = new FieldMirror;
  m->host = fa; m->grid = mirror( fa->g );
  const GridDev & d = m->grid->dev;
  VPB_CUDA( cudaMalloc( &m->d, sizeof(float)*F_NCOMP*(size_t)d.nvp ) );
  VPB_CUDA( cudaMemsetAsync( m->d, 0, sizeof(float)*F_NCOMP*(size_t)d.nvp, r.stream ) );
  VPB_CUDA( cudaMalloc( &m->d_en, sizeof(double)*64 ) );
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  if( !prm || !prm->mc || prm->n_mc<1 ) VPB_ERROR( "field array has no material coefficient table" );
  m->n_mc = prm->n_mc;
  VPB_CUDA( cudaMalloc( &m->d_mc, sizeof(vpb_material_coefficient_t)*(size_t)m->n_mc ) );
  VPB_CUDA( cudaMemcpy( m->d_mc, prm->mc, sizeof(vpb_material_coefficient_t)*(size_t)m->n_mc, cudaMemcpyHostToDevice ) );
  if( m->n_mc>1 ) VPB_CUDA( cudaMalloc( &m->d_mat, sizeof(int16_t)*8*(size_t)d.nvp ) );
  // packed faces: 2 faces x (send+recv) of the largest plane; the largest message
  // (synchronize_tang_e_norm_b) carries 5 values per plane point
  size_t a = (size_t)(d.nx+2)*(d.ny+2), b = (size_t)(d.ny+2)*(d.nz+2), c = (size_t)(d.nz+2)*(d.nx+2);
  size_t mx = a>b ? a : b; if( c>mx ) mx = c;
  m->face_floats = 5*mx + 32;
  VPB_CUDA( cudaMalloc( &m->d_face, sizeof(float)*12*m->face_floats ) );   // [send_lo|send_hi|recv_lo|recv_hi] x 3 axes
  m->dev_valid = false; m->host_valid = true;
  r.fields[fa] = m;
  return m;
}

void to_device( FieldMirror * m ) {
  Runtime & r = rt();
  if( m->dev_valid && r.coherence!=VPB_COHERENCE_STRICT ) return;
  if( m->dev_valid && !m->host_valid ) return;     // device is the only valid copy
  if( !m->host->f ) { if( m->dev_valid ) return; VPB_ERROR( "field array has neither host nor device data" ); }
  const GridDev & d = m->grid->dev;
  if( m->dev_valid && track_clean( m->host->f, sizeof(vpb_field_t)*(size_t)d.nv ) ) return;      // TRACKED: the host did not write it
  if( !m->d_stage ) VPB_CUDA( cudaMalloc( &m->d_stage, sizeof(vpb_field_t)*(size_t)d.nv ) );
  pin_host( m->host->f, sizeof(vpb_field_t)*(size_t)d.nv );
  h2d( m->d_stage, m->host->f, sizeof(vpb_field_t)*(size_t)d.nv, r.stream );
  k_field_aos_to_soa<<< div_up(d.nv,256), 256, 0, r.stream >>>( (const vpb_field_t *)m->d_stage, m->d, m->d_mat, d.nv, d.nvp );
  VPB_CHECK_LAUNCH(); r.launches++;
  m->dev_valid = true; m->stage_has_ids = true;
  track_arm( m->host->f, sizeof(vpb_field_t)*(size_t)d.nv );
}

void to_host( FieldMirror * m ) {
  Runtime & r = rt();
  if( !m->dev_valid || m->host_valid ) return;
  if( !m->host->f ) return;                         // device-only array
  const GridDev & d = m->grid->dev;
  if( !m->d_stage ) VPB_CUDA( cudaMalloc( &m->d_stage, sizeof(vpb_field_t)*(size_t)d.nv ) );
  // keep the material ids of the host copy: the stage still holds the bytes of
  // the last upload (only the 16 floats per voxel are rewritten below)
  pin_host( m->host->f, sizeof(vpb_field_t)*(size_t)d.nv );
  if( !m->stage_has_ids ) {
    h2d( m->d_stage, m->host->f, sizeof(vpb_field_t)*(size_t)d.nv, r.stream );
    m->stage_has_ids = true;
  }
  k_field_soa_to_aos<<< div_up(d.nv,256), 256, 0, r.stream >>>( (vpb_field_t *)m->d_stage, m->d, d.nv, d.nvp );
  VPB_CHECK_LAUNCH(); r.launches++;
  d2h( m->host->f, m->d_stage, sizeof(vpb_field_t)*(size_t)d.nv, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  m->host_valid = true;
  track_arm( m->host->f, sizeof(vpb_field_t)*(size_t)d.nv );
}

} // namespace vpb
