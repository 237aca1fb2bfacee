// TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
//
// Tiny helper linked into oracle/_ref/libvpic_ref_*.so next to the unmodified
// reference objects.  It only (a) boots the reference's services with a given
// thread count so that its C entry points can be driven from ctypes, and
// (b) reports the reference's own sizeof/offsetof for every POD on the hot
// path, which tests/test_abi.py compares with include/vpic_b200/abi.h.
#include "src/vpic/vpic.h"
#include <cstddef>
#include <cstdio>
#include <cstring>

extern "C" {

int refshim_boot( int tpp ) {
  static int booted = 0;
  if( booted ) return thread.n_pipeline;
  static char a0[] = "refshim", a1[] = "--tpp", a2[16];
  snprintf( a2, sizeof(a2), "%d", tpp>0 ? tpp : 1 );
  static char * argv_store[4] = { a0, a1, a2, NULL };
  char ** argv = argv_store;
  int argc = 3;
  boot_services( &argc, &argv );
  booted = 1;
  return thread.n_pipeline;
}

int refshim_n_pipeline( void ) { return thread.n_pipeline; }

// name -> value table of the layout facts the product relies on.
#define FACT(name,val) { name, (long)(val) }
struct refshim_fact { const char * name; long value; };

static const refshim_fact facts[] = {
  FACT( "sizeof(particle_t)",          sizeof(particle_t) ),
  FACT( "sizeof(particle_mover_t)",    sizeof(particle_mover_t) ),
  FACT( "sizeof(particle_injector_t)", sizeof(particle_injector_t) ),
  FACT( "sizeof(species_t)",           sizeof(species_t) ),
  FACT( "species_t.name",              offsetof(species_t,name) ),
  FACT( "species_t.q",                 offsetof(species_t,q) ),
  FACT( "species_t.m",                 offsetof(species_t,m) ),
  FACT( "species_t.np",                offsetof(species_t,np) ),
  FACT( "species_t.max_np",            offsetof(species_t,max_np) ),
  FACT( "species_t.p",                 offsetof(species_t,p) ),
  FACT( "species_t.nm",                offsetof(species_t,nm) ),
  FACT( "species_t.max_nm",            offsetof(species_t,max_nm) ),
  FACT( "species_t.pm",                offsetof(species_t,pm) ),
  FACT( "species_t.last_sorted",       offsetof(species_t,last_sorted) ),
  FACT( "species_t.sort_interval",     offsetof(species_t,sort_interval) ),
  FACT( "species_t.sort_out_of_place", offsetof(species_t,sort_out_of_place) ),
  FACT( "species_t.partition",         offsetof(species_t,partition) ),
  FACT( "species_t.g",                 offsetof(species_t,g) ),
  FACT( "species_t.id",                offsetof(species_t,id) ),
  FACT( "species_t.next",              offsetof(species_t,next) ),
  FACT( "sizeof(interpolator_t)",      sizeof(interpolator_t) ),
  FACT( "sizeof(interpolator_array_t)",sizeof(interpolator_array_t) ),
  FACT( "interpolator_array_t.i",      offsetof(interpolator_array_t,i) ),
  FACT( "interpolator_array_t.g",      offsetof(interpolator_array_t,g) ),
  FACT( "sizeof(accumulator_t)",       sizeof(accumulator_t) ),
  FACT( "sizeof(accumulator_array_t)", sizeof(accumulator_array_t) ),
  FACT( "accumulator_array_t.a",       offsetof(accumulator_array_t,a) ),
  FACT( "accumulator_array_t.n_pipeline", offsetof(accumulator_array_t,n_pipeline) ),
  FACT( "accumulator_array_t.stride",  offsetof(accumulator_array_t,stride) ),
  FACT( "accumulator_array_t.g",       offsetof(accumulator_array_t,g) ),
  FACT( "sizeof(field_t)",             sizeof(field_t) ),
  FACT( "field_t.ex",                  offsetof(field_t,ex) ),
  FACT( "field_t.div_e_err",           offsetof(field_t,div_e_err) ),
  FACT( "field_t.cbx",                 offsetof(field_t,cbx) ),
  FACT( "field_t.div_b_err",           offsetof(field_t,div_b_err) ),
  FACT( "field_t.tcax",                offsetof(field_t,tcax) ),
  FACT( "field_t.rhob",                offsetof(field_t,rhob) ),
  FACT( "field_t.jfx",                 offsetof(field_t,jfx) ),
  FACT( "field_t.rhof",                offsetof(field_t,rhof) ),
  FACT( "field_t.ematx",               offsetof(field_t,ematx) ),
  FACT( "field_t.nmat",                offsetof(field_t,nmat) ),
  FACT( "field_t.fmatx",               offsetof(field_t,fmatx) ),
  FACT( "field_t.cmat",                offsetof(field_t,cmat) ),
  FACT( "sizeof(field_advance_kernels_t)", sizeof(field_advance_kernels_t) ),
  FACT( "sizeof(field_array_t)",       sizeof(field_array_t) ),
  FACT( "field_array_t.f",             offsetof(field_array_t,f) ),
  FACT( "field_array_t.g",             offsetof(field_array_t,g) ),
  FACT( "field_array_t.params",        offsetof(field_array_t,params) ),
  FACT( "field_array_t.kernel",        offsetof(field_array_t,kernel) ),
  FACT( "sizeof(grid_t)",              sizeof(grid_t) ),
  FACT( "grid_t.dt",                   offsetof(grid_t,dt) ),
  FACT( "grid_t.cvac",                 offsetof(grid_t,cvac) ),
  FACT( "grid_t.eps0",                 offsetof(grid_t,eps0) ),
  FACT( "grid_t.step",                 offsetof(grid_t,step) ),
  FACT( "grid_t.t0",                   offsetof(grid_t,t0) ),
  FACT( "grid_t.x0",                   offsetof(grid_t,x0) ),
  FACT( "grid_t.x1",                   offsetof(grid_t,x1) ),
  FACT( "grid_t.nx",                   offsetof(grid_t,nx) ),
  FACT( "grid_t.dx",                   offsetof(grid_t,dx) ),
  FACT( "grid_t.dV",                   offsetof(grid_t,dV) ),
  FACT( "grid_t.rdx",                  offsetof(grid_t,rdx) ),
  FACT( "grid_t.r8V",                  offsetof(grid_t,r8V) ),
  FACT( "grid_t.sx",                   offsetof(grid_t,sx) ),
  FACT( "grid_t.nv",                   offsetof(grid_t,nv) ),
  FACT( "grid_t.bc",                   offsetof(grid_t,bc) ),
  FACT( "grid_t.range",                offsetof(grid_t,range) ),
  FACT( "grid_t.neighbor",             offsetof(grid_t,neighbor) ),
  FACT( "grid_t.rangel",               offsetof(grid_t,rangel) ),
  FACT( "grid_t.rangeh",               offsetof(grid_t,rangeh) ),
  FACT( "grid_t.mp",                   offsetof(grid_t,mp) ),
  FACT( "sizeof(material_coefficient_t)", 64 ),
};

int refshim_n_facts( void ) { return (int)(sizeof(facts)/sizeof(facts[0])); }
const char * refshim_fact_name( int i ) { return facts[i].name; }
long refshim_fact_value( int i ) { return facts[i].value; }

} // extern "C"
