// vpb_host.cu -- stand-alone host layer (vpic_b200.h part 3): our own
// constructors for the POD structs of abi.h, mirroring the reference's
//   src/grid/{grid_structors.c,ops.c:18-211,partition.c:25-182}
//   src/species_advance/species_advance.c:102-146
//   src/sf_interface/{interpolator_array.cc,accumulator_array.c:52-73}
//   src/field_advance/standard/sfa.c:55-216, src/material/material.c
// so that programs which do not link the reference (bench.py, the tests) can
// build exactly the structures the drop-in entry points consume.
#include "vpb_runtime.h"
#include <cstring>
#include <cmath>
#include <climits>

// The reference's checkpoint service (src/util/checkpt/checkpt.h:150-260).  A
// host program built from the reference carries these symbols and expects every
// object its simulation points at to be registered with them (a deck's
// checkpt() call fails on "a pointer to an unregistered object" otherwise);
// they are weak so that the library also loads on its own.
extern "C" {
typedef void   (*vpb_checkpt_func_t)( const void * obj );
typedef void * (*vpb_restore_func_t)( void );
typedef void   (*vpb_reanimate_func_t)( void * obj );
void   register_object( void * obj, vpb_checkpt_func_t, vpb_restore_func_t, vpb_reanimate_func_t ) __attribute__((weak));
void   unregister_object( void * obj ) __attribute__((weak));
void   checkpt_data( const void * data, size_t sz_ele, size_t str_ele, size_t n_ele, size_t max_ele, size_t align ) __attribute__((weak));
void * restore_data( void ) __attribute__((weak));
void   checkpt_ptr( const void * ptr ) __attribute__((weak));
void * restore_ptr( void ) __attribute__((weak));
void   checkpt_sym( const void * saddr ) __attribute__((weak));
void * restore_sym( void ) __attribute__((weak));
// ... and its aligned allocator (src/util/util_base.h:364-387): arrays the
// host program allocated are resized with the host program's own functions.
void   util_malloc_aligned( const char * err, void * mem_ref, size_t n, size_t a ) __attribute__((weak));
void   util_free_aligned( void * mem_ref ) __attribute__((weak));
}
#include <set>

using namespace vpb;

namespace {

static std::set<const void *> & own_species() { static std::set<const void *> s; return s; }
// field arrays that came out of vpic_b200_restore_fa: their f / mc arrays were
// allocated by the host program's restore_data (MALLOC_ALIGNED), not by us
static std::set<const void *> & restored_fas() { static std::set<const void *> s; return s; }

void * alloc_aligned( size_t bytes, size_t align ) {
  void * p = nullptr;
  if( bytes==0 ) bytes = align;
  if( posix_memalign( &p, align, bytes ) ) VPB_ERROR( "Failed to allocate %zu bytes", bytes );
  return p;
}

} // namespace

namespace vpb {

// boundary_p.cc:346-394 for HOST-resident particle / mover arrays: a new array
// of n elements, the old contents copied over, the old array released.  The
// arrays of a species made by vpic_b200_new_species are this library's; any
// other species belongs to the host program and is resized with the host
// program's allocator (the reference's MALLOC_ALIGNED / FREE_ALIGNED).
static void * host_array_resize( vpb_species_t * sp, void * old, size_t old_bytes, size_t new_bytes, size_t keep_bytes ) {
  const bool ours = own_species().count( sp )!=0;
  void * np_ = nullptr;
  if( ours ) np_ = alloc_aligned( new_bytes, 128 );
  else if( util_malloc_aligned && util_free_aligned )
    util_malloc_aligned( "libvpic_b200: resizing storage of a species failed (%lu bytes, %lu aligned)", &np_, new_bytes, 128 );
  else VPB_ERROR( "species \"%s\": its arrays were not allocated by this library and the host program has no "
                  "util_malloc_aligned / util_free_aligned to resize them with; raise max_np / max_nm", sp->name );
  // (keep_bytes: how much of the old array means anything -- the elements in use, or nothing at all when the
  //  device copy is the current one: RESIDENT runs do not copy gigabytes of stale particles around)
  size_t n = old_bytes<new_bytes ? old_bytes : new_bytes;
  if( keep_bytes<n ) n = keep_bytes;
  if( n ) memcpy( np_, old, n );
  unpin_host( old );
  if( ours ) free( old ); else { void * o = old; util_free_aligned( &o ); }
  return np_;
}
void host_resize_particles( vpb_species_t * sp, size_t n, size_t keep ) {
  sp->p = (vpb_particle_t *)host_array_resize( sp, sp->p, sizeof(vpb_particle_t)*(size_t)sp->max_np, sizeof(vpb_particle_t)*n, sizeof(vpb_particle_t)*keep );
  sp->max_np = (int)n;
}
void host_resize_movers( vpb_species_t * sp, size_t n, size_t keep ) {
  sp->pm = (vpb_particle_mover_t *)host_array_resize( sp, sp->pm, sizeof(vpb_particle_mover_t)*(size_t)sp->max_nm, sizeof(vpb_particle_mover_t)*n, sizeof(vpb_particle_mover_t)*keep );
  sp->max_nm = (int)n;
}

} // namespace vpb

namespace {

#define LOCAL_CELL_ID(x,y,z) VPB_VOXEL(x,y,z,lnx,lny,lnz)

// ops.c:18-86
void size_grid( vpb_grid_t * g, int lnx, int lny, int lnz ) {
  if( !g || lnx<1 || lny<1 || lnz<1 ) VPB_ERROR( "Bad args" );
  g->sx = 1; g->sy = (lnx+2)*g->sx; g->sz = (lny+2)*g->sy; g->nv = (lnz+2)*g->sz;
  g->nx = lnx; g->ny = lny; g->nz = lnz;
  for( int k=-1; k<=1; k++ ) for( int j=-1; j<=1; j++ ) for( int i=-1; i<=1; i++ )
    g->bc[ VPB_BOUNDARY(i,j,k) ] = VPB_PEC_FIELDS;
  g->bc[ VPB_BOUNDARY(0,0,0) ] = g_world_rank;
  free( g->range );
  g->range = (int64_t *)alloc_aligned( sizeof(int64_t)*(size_t)(g_world_size+1), 16 );
  // every rank of a uniform box partition owns the same number of voxels,
  // which is what the reference's allgather (ops.c:44) would report
  for( int i=0; i<=g_world_size; i++ ) g->range[i] = (int64_t)i*g->nv;
  g->rangel = g->range[g_world_rank];
  g->rangeh = g->range[g_world_rank+1]-1;
  free( g->neighbor );
  g->neighbor = (int64_t *)alloc_aligned( sizeof(int64_t)*6*(size_t)g->nv, 128 );
  for( int z=0; z<=lnz+1; z++ ) for( int y=0; y<=lny+1; y++ ) for( int x=0; x<=lnx+1; x++ ) {
    int64_t * nb = g->neighbor + 6*(int64_t)LOCAL_CELL_ID(x,y,z);
    const bool ghost = x==0 || x==lnx+1 || y==0 || y==lny+1 || z==0 || z==lnz+1;
    nb[0] = ( ghost || x==1   ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x-1,y,z);
    nb[1] = ( ghost || y==1   ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x,y-1,z);
    nb[2] = ( ghost || z==1   ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x,y,z-1);
    nb[3] = ( ghost || x==lnx ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x+1,y,z);
    nb[4] = ( ghost || y==lny ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x,y+1,z);
    nb[5] = ( ghost || z==lnz ) ? VPB_REFLECT_PARTICLES : g->rangel + LOCAL_CELL_ID(x,y,z+1);
  }
}

// ops.c:118-163: glue one face to the same-shaped brick owned by `rank`.
void join_grid( vpb_grid_t * g, int boundary, int rank ) {
  if( !g || boundary<0 || boundary>=27 || boundary==VPB_BOUNDARY(0,0,0) || rank<0 || rank>=g_world_size )
    VPB_ERROR( "Bad args" );
  g->bc[boundary] = rank;
  const int lnx = g->nx, lny = g->ny, lnz = g->nz;
  const int ln[3] = { lnx, lny, lnz };
  const int64_t rnc = g->range[rank+1] - g->range[rank];
  static const int bnd[6] = { VPB_BOUNDARY(-1,0,0), VPB_BOUNDARY(0,-1,0), VPB_BOUNDARY(0,0,-1),
                              VPB_BOUNDARY( 1,0,0), VPB_BOUNDARY(0, 1,0), VPB_BOUNDARY(0,0, 1) };
  for( int tag=0; tag<6; tag++ ) {
    if( boundary!=bnd[tag] ) continue;
    const int X = tag%3, Y = (X+1)%3, Z = (X+2)%3, side = tag<3 ? -1 : 1;
    if( rnc%( (int64_t)(ln[Y]+2)*(ln[Z]+2) )!=0 ) VPB_ERROR( "Remote face is incompatible" );
    int rn[3]; rn[X] = (int)( rnc/( (int64_t)(ln[Y]+2)*(ln[Z]+2) ) ) - 2; rn[Y] = ln[Y]; rn[Z] = ln[Z];
    for( int lz=1; lz<=ln[Z]; lz++ ) for( int ly=1; ly<=ln[Y]; ly++ ) {
      int l[3], r[3];
      l[X] = side<0 ? 1 : ln[X]; l[Y] = ly; l[Z] = lz;
      r[X] = side<0 ? rn[X] : 1; r[Y] = ly; r[Z] = lz;
      g->neighbor[ 6*(int64_t)VPB_VOXEL(l[0],l[1],l[2],lnx,lny,lnz) + tag ] =
        g->range[rank] + VPB_VOXEL(r[0],r[1],r[2],rn[0],rn[1],rn[2]);
    }
    return;
  }
}

void rank_to_index( int rank, int gpx, int gpy, int & ix, int & iy, int & iz ) {
  ix = rank; iy = ix/gpx; ix -= iy*gpx; iz = iy/gpy; iy -= iz*gpy;
}
int index_to_rank( int ix, int iy, int iz, int gpx, int gpy, int gpz ) {
  while( ix>=gpx ) ix -= gpx; while( ix<0 ) ix += gpx;
  while( iy>=gpy ) iy -= gpy; while( iy<0 ) iy += gpy;
  while( iz>=gpz ) iz -= gpz; while( iz<0 ) iz += gpz;
  return ix + gpx*( iy + gpy*iz );
}

} // namespace

extern "C" {

vpb_grid_t * vpic_b200_new_grid( void ) {
  vpb_grid_t * g = (vpb_grid_t *)calloc( 1, sizeof(vpb_grid_t) );
  if( !g ) VPB_ERROR( "Failed to allocate grid" );
  for( int i=0; i<27; i++ ) g->bc[i] = VPB_PEC_FIELDS;
  g->bc[ VPB_BOUNDARY(0,0,0) ] = g_world_rank;
  return g;
}

void vpic_b200_delete_grid( vpb_grid_t * g ) {
  if( !g ) return;
  vpic_b200_forget( g );
  free( g->neighbor ); free( g->range ); free( g );
}

// partition.c:25-80
void vpic_b200_partition_periodic_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                       double gx1, double gy1, double gz1,
                                       int gnx, int gny, int gnz, int gpx, int gpy, int gpz ) {
  if( !g ) VPB_ERROR( "NULL grid" );
  if( gpx<1 || gpy<1 || gpz<1 || gpx*gpy*gpz!=g_world_size )
    VPB_ERROR( "Bad domain decompostion (%ix%ix%i)", gpx, gpy, gpz );
  if( gnx<1 || gny<1 || gnz<1 || gnx%gpx!=0 || gny%gpy!=0 || gnz%gpz!=0 )
    VPB_ERROR( "Bad resolution (%ix%ix%i) for domain decomposition (%ix%ix%i)", gnx, gny, gnz, gpx, gpy, gpz );
  int px, py, pz; rank_to_index( g_world_rank, gpx, gpy, px, py, pz );
  g->dx = (gx1-gx0)/(double)gnx; g->dy = (gy1-gy0)/(double)gny; g->dz = (gz1-gz0)/(double)gnz;
  g->dV = ((gx1-gx0)/(double)gnx)*((gy1-gy0)/(double)gny)*((gz1-gz0)/(double)gnz);
  g->rdx = (double)gnx/(gx1-gx0); g->rdy = (double)gny/(gy1-gy0); g->rdz = (double)gnz/(gz1-gz0);
  g->r8V = ((double)gnx/(gx1-gx0))*((double)gny/(gy1-gy0))*((double)gnz/(gz1-gz0))*0.125;
  double f;
  f = (double) px   /(double)gpx; g->x0 = gx0*(1-f) + gx1*f;
  f = (double) py   /(double)gpy; g->y0 = gy0*(1-f) + gy1*f;
  f = (double) pz   /(double)gpz; g->z0 = gz0*(1-f) + gz1*f;
  f = (double)(px+1)/(double)gpx; g->x1 = gx0*(1-f) + gx1*f;
  f = (double)(py+1)/(double)gpy; g->y1 = gy0*(1-f) + gy1*f;
  f = (double)(pz+1)/(double)gpz; g->z1 = gz0*(1-f) + gz1*f;
  size_grid( g, gnx/gpx, gny/gpy, gnz/gpz );
  join_grid( g, VPB_BOUNDARY(-1, 0, 0), index_to_rank( px-1, py,   pz,   gpx, gpy, gpz ) );
  join_grid( g, VPB_BOUNDARY( 0,-1, 0), index_to_rank( px,   py-1, pz,   gpx, gpy, gpz ) );
  join_grid( g, VPB_BOUNDARY( 0, 0,-1), index_to_rank( px,   py,   pz-1, gpx, gpy, gpz ) );
  join_grid( g, VPB_BOUNDARY( 1, 0, 0), index_to_rank( px+1, py,   pz,   gpx, gpy, gpz ) );
  join_grid( g, VPB_BOUNDARY( 0, 1, 0), index_to_rank( px,   py+1, pz,   gpx, gpy, gpz ) );
  join_grid( g, VPB_BOUNDARY( 0, 0, 1), index_to_rank( px,   py,   pz+1, gpx, gpy, gpz ) );
  vpic_b200_forget( g );   // any cached device view of this grid is stale now
}

// ops.c:165-211
void vpic_b200_set_fbc( vpb_grid_t * g, int boundary, int fbc ) {
  if( !g || boundary<0 || boundary>=27 || boundary==VPB_BOUNDARY(0,0,0) ||
      ( fbc!=VPB_PEC_FIELDS && fbc!=VPB_SYMMETRIC_FIELDS && fbc!=VPB_PMC_FIELDS && fbc!=VPB_ABSORB_FIELDS ) )
    VPB_ERROR( "Bad args" );
  g->bc[boundary] = fbc;
  vpic_b200_grid_changed( g );
}

void vpic_b200_set_pbc( vpb_grid_t * g, int boundary, int pbc ) {
  if( !g || boundary<0 || boundary>=27 || boundary==VPB_BOUNDARY(0,0,0) || pbc>=0 ) VPB_ERROR( "Bad args" );
  const int lnx = g->nx, lny = g->ny, lnz = g->nz; const int ln[3] = { lnx, lny, lnz };
  static const int bnd[6] = { VPB_BOUNDARY(-1,0,0), VPB_BOUNDARY(0,-1,0), VPB_BOUNDARY(0,0,-1),
                              VPB_BOUNDARY( 1,0,0), VPB_BOUNDARY(0, 1,0), VPB_BOUNDARY(0,0, 1) };
  for( int tag=0; tag<6; tag++ ) {
    if( boundary!=bnd[tag] ) continue;
    const int X = tag%3, Y = (X+1)%3, Z = (X+2)%3;
    for( int lz=1; lz<=ln[Z]; lz++ ) for( int ly=1; ly<=ln[Y]; ly++ ) {
      int l[3]; l[X] = tag<3 ? 1 : ln[X]; l[Y] = ly; l[Z] = lz;
      g->neighbor[ 6*(int64_t)VPB_VOXEL(l[0],l[1],l[2],lnx,lny,lnz) + tag ] = pbc;
    }
  }
  vpic_b200_grid_changed( g );
}

// partition.c:82-135 and :137-182: periodic box, then override the faces on the
// edge of the global domain.
static void override_edges( vpb_grid_t * g, int gnx, int gny, int gnz, int gpx, int gpy, int gpz, int fbc, int pbc ) {
  int px, py, pz; rank_to_index( g_world_rank, gpx, gpy, px, py, pz );
  struct { bool edge; int b; } e[6] = {
    { px==0     && gnx>1, VPB_BOUNDARY(-1,0,0) }, { py==0     && gny>1, VPB_BOUNDARY(0,-1,0) }, { pz==0     && gnz>1, VPB_BOUNDARY(0,0,-1) },
    { px==gpx-1 && gnx>1, VPB_BOUNDARY( 1,0,0) }, { py==gpy-1 && gny>1, VPB_BOUNDARY(0, 1,0) }, { pz==gpz-1 && gnz>1, VPB_BOUNDARY(0,0, 1) } };
  for( int k=0; k<6; k++ ) if( e[k].edge ) { vpic_b200_set_fbc( g, e[k].b, fbc ); vpic_b200_set_pbc( g, e[k].b, pbc ); }
}

void vpic_b200_partition_absorbing_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                        double gx1, double gy1, double gz1,
                                        int gnx, int gny, int gnz, int gpx, int gpy, int gpz, int pbc ) {
  vpic_b200_partition_periodic_box( g, gx0, gy0, gz0, gx1, gy1, gz1, gnx, gny, gnz, gpx, gpy, gpz );
  override_edges( g, gnx, gny, gnz, gpx, gpy, gpz, VPB_ABSORB_FIELDS, pbc );
}

void vpic_b200_partition_metal_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                    double gx1, double gy1, double gz1,
                                    int gnx, int gny, int gnz, int gpx, int gpy, int gpz ) {
  vpic_b200_partition_periodic_box( g, gx0, gy0, gz0, gx1, gy1, gz1, gnx, gny, gnz, gpx, gpy, gpz );
  override_edges( g, gnx, gny, gnz, gpx, gpy, gpz, VPB_PEC_FIELDS, VPB_REFLECT_PARTICLES );
}

// species_advance.c:102-146
vpb_species_t * vpic_b200_new_species( const char * name, float q, float m, size_t max_np, size_t max_nm,
                                       int sort_interval, int sort_out_of_place, vpb_grid_t * g, int host_arrays ) {
  const size_t len = name ? strlen( name ) : 0;
  if( !len ) VPB_ERROR( "Cannot create a nameless species" );
  if( !g ) VPB_ERROR( "NULL grid" );
  if( g->nv==0 ) VPB_ERROR( "Allocate grid before defining species." );
  if( max_np<1 ) max_np = 1;
  if( max_nm<1 ) max_nm = 1;
  if( max_np>(size_t)INT_MAX || max_nm>(size_t)INT_MAX ) VPB_ERROR( "species_t counts are 32-bit (species_advance_aos.h:59)" );
  vpb_species_t * sp = (vpb_species_t *)calloc( 1, sizeof(vpb_species_t) );
  sp->name = strdup( name );
  sp->q = q; sp->m = m;
  sp->max_np = (int)max_np; sp->max_nm = (int)max_nm;
  if( host_arrays ) {
    sp->p  = (vpb_particle_t *)alloc_aligned( sizeof(vpb_particle_t)*max_np, 128 );
    sp->pm = (vpb_particle_mover_t *)alloc_aligned( sizeof(vpb_particle_mover_t)*max_nm, 128 );
  }
  sp->last_sorted = INT64_MIN;
  sp->sort_interval = sort_interval; sp->sort_out_of_place = sort_out_of_place;
  sp->partition = (int *)alloc_aligned( sizeof(int)*( (size_t)g->nv + 1 ), 128 );
  sp->g = g;
  own_species().insert( sp );
  return sp;
}

void vpic_b200_delete_species( vpb_species_t * sp ) {
  if( !sp ) return;
  own_species().erase( sp );
  vpic_b200_forget( sp );
  free( sp->partition ); free( sp->pm ); free( sp->p ); free( sp->name ); free( sp );
}

vpb_interpolator_array_t * vpic_b200_new_interpolator_array( vpb_grid_t * g ) {
  if( !g ) VPB_ERROR( "NULL grid" );
  vpb_interpolator_array_t * ia = (vpb_interpolator_array_t *)calloc( 1, sizeof(*ia) );
  ia->i = (vpb_interpolator_t *)alloc_aligned( sizeof(vpb_interpolator_t)*(size_t)g->nv, 128 );
  memset( ia->i, 0, sizeof(vpb_interpolator_t)*(size_t)g->nv );
  ia->g = g;
  return ia;
}
void vpic_b200_delete_interpolator_array( vpb_interpolator_array_t * ia ) {
  if( !ia ) return;
  vpic_b200_forget( ia ); free( ia->i ); free( ia );
}

vpb_accumulator_array_t * vpic_b200_new_accumulator_array( vpb_grid_t * g, int n_pipeline ) {
  if( !g ) VPB_ERROR( "Bad grid." );
  if( n_pipeline<0 ) n_pipeline = 0;
  vpb_accumulator_array_t * aa = (vpb_accumulator_array_t *)calloc( 1, sizeof(*aa) );
  aa->n_pipeline = n_pipeline;
  aa->stride = ( g->nv + 1 ) & ~1;           // POW2_CEIL(nv,2)
  aa->g = g;
  const size_t n = (size_t)(n_pipeline+1)*(size_t)aa->stride;
  aa->a = (vpb_accumulator_t *)alloc_aligned( sizeof(vpb_accumulator_t)*n, 128 );
  memset( aa->a, 0, sizeof(vpb_accumulator_t)*n );
  return aa;
}
void vpic_b200_delete_accumulator_array( vpb_accumulator_array_t * aa ) {
  if( !aa ) return;
  vpic_b200_forget( aa ); free( aa->a ); free( aa );
}

vpb_material_t * vpic_b200_new_material( const char * name, float eps, float mu, float sigma ) {
  vpb_material_t * m = (vpb_material_t *)calloc( 1, sizeof(*m) );
  m->name = strdup( name ? name : "material" );
  m->epsx = m->epsy = m->epsz = eps;
  m->mux = m->muy = m->muz = mu;
  m->sigmax = m->sigmay = m->sigmaz = sigma;
  m->id = 0; m->next = nullptr;
  return m;
}

// material.c:69-86: ids count up from 0, the newest material is the list head
vpb_material_t * vpic_b200_append_material( vpb_material_t * m, vpb_material_t ** m_list ) {
  if( !m || !m_list ) VPB_ERROR( "Bad args" );
  if( m->next ) VPB_ERROR( "Material \"%s\" already in a list", m->name );
  for( const vpb_material_t * k=*m_list; k; k=k->next )
    if( strcmp( k->name, m->name )==0 ) VPB_ERROR( "There is already a material named \"%s\" in list", m->name );
  const int id = *m_list ? (*m_list)->id+1 : 0;
  if( id>=32767 ) VPB_ERROR( "Too many materials in list to append material \"%s\"", m->name );
  m->id = (vpb_material_id)id;
  m->next = *m_list;
  *m_list = m;
  return m;
}

// sfa.c:55-151 (coefficients) and :191-216 (constructor)
vpb_field_array_t * new_standard_field_array( vpb_grid_t * g, const vpb_material_t * m_list, float damp ) {
  if( !g || !m_list || damp<0 ) VPB_ERROR( "Bad args" );
  vpb_field_array_t * fa = (vpb_field_array_t *)calloc( 1, sizeof(*fa) );
  fa->f = (vpb_field_t *)alloc_aligned( sizeof(vpb_field_t)*(size_t)g->nv, 128 );
  memset( fa->f, 0, sizeof(vpb_field_t)*(size_t)g->nv );
  fa->g = g;
  int n_mc = 0;
  for( const vpb_material_t * m=m_list; m; m=m->next ) n_mc++;
  vpb_sfa_params_t * p = (vpb_sfa_params_t *)calloc( 1, sizeof(*p) );
  p->mc = (vpb_material_coefficient_t *)alloc_aligned( sizeof(vpb_material_coefficient_t)*(size_t)(n_mc+2), 128 );
  memset( p->mc, 0, sizeof(vpb_material_coefficient_t)*(size_t)(n_mc+2) );
  p->n_mc = n_mc; p->damp = damp;
  for( const vpb_material_t * m=m_list; m; m=m->next ) {
    vpb_material_coefficient_t * mc = p->mc + m->id;
    float ax = ( m->sigmax*g->dt )/( m->epsx*g->eps0 );
    float ay = ( m->sigmay*g->dt )/( m->epsy*g->eps0 );
    float az = ( m->sigmaz*g->dt )/( m->epsz*g->eps0 );
    mc->decayx = exp(-ax); mc->decayy = exp(-ay); mc->decayz = exp(-az);
    if( ax==0 ) mc->drivex = 1./m->epsx; else if( mc->decayx==0 ) mc->drivex = 0;
    else mc->drivex = 2.*exp(-0.5*ax)*sinh(0.5*ax)/(ax*m->epsx);
    if( ay==0 ) mc->drivey = 1./m->epsy; else if( mc->decayy==0 ) mc->drivey = 0;
    else mc->drivey = 2.*exp(-0.5*ay)*sinh(0.5*ay)/(ay*m->epsy);
    if( az==0 ) mc->drivez = 1./m->epsz; else if( mc->decayz==0 ) mc->drivez = 0;
    else mc->drivez = 2.*exp(-0.5*az)*sinh(0.5*az)/(az*m->epsz);
    mc->rmux = 1./m->mux; mc->rmuy = 1./m->muy; mc->rmuz = 1./m->muz;
    mc->nonconductive = ( ax==0 && ay==0 && az==0 ) ? 1. : 0.;
    mc->epsx = m->epsx; mc->epsy = m->epsy; mc->epsz = m->epsz;
  }
  fa->params = p;
  memset( fa->kernel, 0, sizeof(fa->kernel) );
  if( register_object && checkpt_data && checkpt_ptr && checkpt_sym )
    register_object( fa, vpic_b200_checkpt_fa, vpic_b200_restore_fa, nullptr );
  fa->kernel->delete_fa      = vpic_b200_delete_fa;
  fa->kernel->advance_b      = vpic_b200_advance_b;
  fa->kernel->advance_e      = vpic_b200_advance_e;
  fa->kernel->energy_f       = vpic_b200_energy_f;
  fa->kernel->clear_jf       = vpic_b200_clear_jf;
  fa->kernel->synchronize_jf = vpic_b200_synchronize_jf;
  fa->kernel->clear_rhof     = vpic_b200_clear_rhof;
  fa->kernel->synchronize_rho           = vpic_b200_synchronize_rho;
  fa->kernel->compute_rhob              = vpic_b200_compute_rhob;
  fa->kernel->compute_curl_b            = vpic_b200_compute_curl_b;
  fa->kernel->synchronize_tang_e_norm_b = vpic_b200_synchronize_tang_e_norm_b;
  fa->kernel->compute_div_e_err         = vpic_b200_compute_div_e_err;
  fa->kernel->compute_rms_div_e_err     = vpic_b200_compute_rms_div_e_err;
  fa->kernel->clean_div_e               = vpic_b200_clean_div_e;
  fa->kernel->compute_div_b_err         = vpic_b200_compute_div_b_err;
  fa->kernel->compute_rms_div_b_err     = vpic_b200_compute_rms_div_b_err;
  fa->kernel->clean_div_b               = vpic_b200_clean_div_b;
  return fa;
}

// sfa.c:161-188 (checkpt_standard_field_array / restore_standard_field_array)
// with field_advance.c:17-58 for the kernel table; the host copy is made
// current first, the device mirror of a restored array is rebuilt on first use.
void vpic_b200_checkpt_fa( const void * obj ) {
  const vpb_field_array_t * fa = (const vpb_field_array_t *)obj;
  if( !fa || !checkpt_data ) VPB_ERROR( "Bad args" );
  vpic_b200_sync_to_host( const_cast<vpb_field_array_t *>( fa ) );
  const vpb_sfa_params_t * p = (const vpb_sfa_params_t *)fa->params;
  checkpt_data( fa, sizeof(*fa), sizeof(*fa), 1, 1, 0 );
  { const size_t sz = sizeof(vpb_field_t)*(size_t)fa->g->nv; checkpt_data( fa->f, sz, sz, 1, 1, 128 ); }
  checkpt_ptr( fa->g );
  checkpt_data( p, sizeof(*p), sizeof(*p), 1, 1, 0 );
  { const size_t sz = sizeof(vpb_material_coefficient_t)*(size_t)p->n_mc; checkpt_data( p->mc, sz, sz, 1, 1, 128 ); }
  void * const * k = (void * const *)fa->kernel;
  for( size_t i=0; i<sizeof(vpb_field_advance_kernels_t)/sizeof(void *); i++ ) checkpt_sym( k[i] );
}

void * vpic_b200_restore_fa( void ) {
  if( !restore_data || !restore_ptr || !restore_sym ) VPB_ERROR( "no checkpoint service in this program" );
  vpb_field_array_t * fa = (vpb_field_array_t *)restore_data();
  fa->f = (vpb_field_t *)restore_data();
  fa->g = (vpb_grid_t *)restore_ptr();
  vpb_sfa_params_t * p = (vpb_sfa_params_t *)restore_data();
  p->mc = (vpb_material_coefficient_t *)restore_data();
  fa->params = p;
  void ** k = (void **)fa->kernel;
  for( size_t i=0; i<sizeof(vpb_field_advance_kernels_t)/sizeof(void *); i++ ) k[i] = restore_sym();
  restored_fas().insert( fa );
  return fa;
}

void vpic_b200_delete_fa( vpb_field_array_t * fa ) {
  if( !fa ) return;
  if( unregister_object && register_object ) unregister_object( fa );
  vpic_b200_forget( fa );
  vpb_sfa_params_t * p = (vpb_sfa_params_t *)fa->params;
  // A restored array's f and mc are MALLOC_ALIGNED blocks of the host program
  // (checkpt.c:424-425: the pointer sits inside a larger malloc block), so they
  // go back through its FREE_ALIGNED; the structs themselves were plain malloc.
  const bool restored = restored_fas().erase( fa )!=0;
  if( restored && !util_free_aligned ) VPB_ERROR( "a restored field array needs the host program's util_free_aligned" );
  auto release = [&]( void * a ) { if( !a ) return; if( restored ) util_free_aligned( &a ); else free( a ); };
  unpin_host( fa->f );
  if( p ) { release( p->mc ); free( p ); }
  release( fa->f ); free( fa );
}

// field_advance.c: delete_field_array dispatches through the table
void delete_field_array( vpb_field_array_t * fa ) {
  if( !fa ) return;
  if( fa->kernel->delete_fa ) fa->kernel->delete_fa( fa ); else vpic_b200_delete_fa( fa );
}

} // extern "C"
