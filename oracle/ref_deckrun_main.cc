// TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
//
// A main() for running an UNMODIFIED reference input deck (any translation
// unit that defines the vpic_simulation::user_* members, e.g. the reference's
// test/unit/energy_comparison/weibel_driver.cc compiled with -Dmain=unused)
// against oracle/_ref/libvpic_ref_scalar.so, dumping the raw host state so
// that tests/golden fixtures can be generated from the reference itself.
//
// Usage: deckrun <out_prefix> [dump_every] [-- deck args...]
//   writes <out_prefix>.init.vpb right after vpic_simulation::initialize()
//   and, if dump_every>0, <out_prefix>.step<N>.vpb every dump_every steps.
//
// File format ("VPBDUMP1"): repeated records
//   char name[48]; uint64 nbytes; uint8 data[nbytes]   (little endian)
// read back by vpic_b200/golden_io.py.
//
// Compile this TU with -Dprotected=public so the driver's state is reachable;
// access specifiers do not change the class layout.
#include "src/vpic/vpic.h"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>

static void put( FILE * f, const char * name, const void * data, uint64_t nbytes ) {
  char nm[48]; memset( nm, 0, sizeof(nm) ); strncpy( nm, name, sizeof(nm)-1 );
  fwrite( nm, 1, sizeof(nm), f );
  fwrite( &nbytes, 8, 1, f );
  if( nbytes ) fwrite( data, 1, nbytes, f );
}

static void dump_state( vpic_simulation & sim, const std::string & fname ) {
  FILE * f = fopen( fname.c_str(), "wb" );
  if( !f ) { fprintf( stderr, "cannot open %s\n", fname.c_str() ); exit(1); }
  fwrite( "VPBDUMP1", 1, 8, f );
  grid_t * g = sim.grid;
  put( f, "grid", g, sizeof(grid_t) );
  put( f, "grid.neighbor", g->neighbor, sizeof(int64_t)*6*(size_t)g->nv );
  put( f, "grid.range", g->range, sizeof(int64_t)*(size_t)(world_size+1) );
  int misc[8] = { sim.num_step, sim.num_comm_round, sim.clean_div_e_interval, sim.clean_div_b_interval,
                  sim.sync_shared_interval, world_rank, world_size, sim.accumulator_array->n_pipeline };
  put( f, "sim.misc", misc, sizeof(misc) );
  put( f, "field.f", sim.field_array->f, sizeof(field_t)*(size_t)g->nv );
  // sfa_params_t { material_coefficient_t * mc; int n_mc; float damp; }
  struct params_view { void * mc; int n_mc; float damp; };
  params_view * pv = (params_view *)sim.field_array->params;
  put( f, "field.mc", pv->mc, 64*(size_t)pv->n_mc );
  put( f, "field.damp", &pv->damp, 4 );
  put( f, "interp.i", sim.interpolator_array->i, sizeof(interpolator_t)*(size_t)g->nv );
  int nsp = 0;
  for( species_t * sp = sim.species_list; sp; sp = sp->next ) nsp++;
  put( f, "species.count", &nsp, 4 );
  int k = 0;
  for( species_t * sp = sim.species_list; sp; sp = sp->next, k++ ) {
    char nm[48];
    snprintf( nm, sizeof(nm), "species%d.name", k );    put( f, nm, sp->name, strlen(sp->name) );
    float qm[2] = { sp->q, sp->m };
    snprintf( nm, sizeof(nm), "species%d.qm", k );      put( f, nm, qm, 8 );
    int ints[6] = { sp->np, sp->max_np, sp->nm, sp->max_nm, sp->sort_interval, sp->sort_out_of_place };
    snprintf( nm, sizeof(nm), "species%d.ints", k );    put( f, nm, ints, sizeof(ints) );
    snprintf( nm, sizeof(nm), "species%d.p", k );       put( f, nm, sp->p, sizeof(particle_t)*(size_t)sp->np );
  }
  fclose( f );
}

int main( int argc, char ** argv ) {
  boot_services( &argc, &argv );
  if( argc<2 ) { fprintf( stderr, "usage: %s out_prefix [dump_every] [deck args]\n", argv[0] ); return 1; }
  std::string prefix = argv[1];
  int dump_every = argc>2 ? atoi( argv[2] ) : 0;
  int dargc = argc>3 ? argc-3 : 0;
  char ** dargv = argc>3 ? argv+3 : NULL;
  {
    vpic_simulation sim;
    sim.initialize( dargc, dargv );
    dump_state( sim, prefix + ".init.vpb" );
    int step = 0;
    while( sim.advance() ) {
      step++;
      if( dump_every>0 && step%dump_every==0 ) {
        char buf[64]; snprintf( buf, sizeof(buf), ".step%d.vpb", step );
        dump_state( sim, prefix + buf );
      }
    }
    step++;
    if( dump_every>0 ) dump_state( sim, prefix + ".final.vpb" );
    sim.finalize();
  }
  halt_services();
  return 0;
}
