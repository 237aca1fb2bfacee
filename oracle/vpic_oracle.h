/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * vpic_oracle: a plain-C, single-threaded CPU restatement of the reference's
 * per-step PIC hot path.  It exists only so that
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg can check the
 * CUDA library; nothing under vpic_b200/ may call it.
 *
 * Semantics are those of the reference's scalar (no USE_V*) build.  The
 * number of "pipelines" whose work split and accumulator copies are emulated
 * is taken from aa->n_pipeline (the reference runs with n_pipeline = --tpp).
 * Floating point: compiled with -ffp-contract=off on baseline x86-64 so every
 * operation is individually rounded to binary32 exactly like the reference's
 * reference-Release build.
 *
 * Parity pin: every function is compared bit-for-bit with the compiled
 * reference (oracle/_ref/libvpic_ref_scalar.so) in tests/test_oracle_vs_ref.py
 * and with the committed golden fixtures in tests/golden/ (generated from the
 * reference by tests/golden/make_golden.py); the 700-step Weibel run must
 * reproduce the reference's own test/unit/energy_comparison/energies_gold.
 */
#ifndef VPIC_ORACLE_H
#define VPIC_ORACLE_H

#include "vpic_b200/abi.h"

#ifdef __cplusplus
extern "C" {
#endif

/* world_rank / world_size as the reference's src/util/mp sees them. */
void orc_set_world( int rank, int size );

/* Ghost-plane transport used by the field exchanges and by boundary_p.  The
   default (NULL) is the single-rank loop-back: a message sent out of face
   (i,j,k) arrives on this rank's own opposite face when g->bc names this
   rank.  Multi-rank tests install a callback that moves the buffers with
   torch.distributed (gloo).
   exchange( user, axis, send_lo, send_hi, recv_lo, recv_hi, nbytes_lo_hi[4] ):
     send_lo goes to the -axis neighbour, send_hi to the +axis neighbour;
     recv_lo is what arrives FROM the -axis neighbour, recv_hi from +axis.
     A NULL send/recv pointer means "no remote neighbour on that side". */
typedef void (*orc_exchange_fn)( void * user, int axis,
                                 const void * send_lo, const void * send_hi,
                                 void * recv_lo, void * recv_hi,
                                 const int64_t * send_bytes /*[2]*/,
                                 int64_t * recv_bytes /*[2] in:capacity out:got*/ );
void orc_set_exchange( orc_exchange_fn fn, void * user );

/* F7  src/sf_interface/pipeline/interpolator_array_pipeline.cc:20-126 */
void orc_load_interpolator_array( vpb_interpolator_array_t * ia, const vpb_field_array_t * fa );
/* F4  src/sf_interface/pipeline/clear_accumulators_pipeline.c:9-61 */
void orc_clear_accumulator_array( vpb_accumulator_array_t * aa );
/* F5  src/sf_interface/pipeline/reduce_accumulators_pipeline.cc:14-243 */
void orc_reduce_accumulator_array( vpb_accumulator_array_t * aa );
/* F6  src/sf_interface/pipeline/unload_accumulator_pipeline.cc:17-137 */
void orc_unload_accumulator_array( vpb_field_array_t * fa, const vpb_accumulator_array_t * aa );

/* F1  src/species_advance/standard/pipeline/advance_p_pipeline.cc:20-314 */
void orc_advance_p( vpb_species_t * sp, vpb_accumulator_array_t * aa, const vpb_interpolator_array_t * ia );
/* F2  src/species_advance/standard/move_p.cc:216-353 */
int  orc_move_p( vpb_particle_t * p0, vpb_particle_mover_t * pm, vpb_accumulator_t * a0,
                 const vpb_grid_t * g, float qsp );
/* F3  src/species_advance/standard/pipeline/sort_p_pipeline.c:143-371 */
void orc_sort_p( vpb_species_t * sp );
/* F13 src/species_advance/standard/pipeline/energy_p_pipeline.cc:17-115 */
double orc_energy_p( const vpb_species_t * sp, const vpb_interpolator_array_t * ia );
/* N2  src/species_advance/standard/pipeline/{un,}center_p_pipeline.cc */
void orc_uncenter_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia );
void orc_center_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia );
/* number of pipelines energy_p splits its double partial sums over (default 1) */
void orc_set_energy_pipelines( int P );

/* F8  src/field_advance/standard/sfa.c:231-237 */
void orc_clear_jf( vpb_field_array_t * fa );
/* F9  src/field_advance/standard/remote.c:417-508 + local.c:335-368 */
void orc_synchronize_jf( vpb_field_array_t * fa );
/* F10 src/field_advance/standard/pipeline/advance_b_pipeline.cc:52-125 + local.c:267-296 */
void orc_advance_b( vpb_field_array_t * fa, float frac );
/* F11 src/field_advance/standard/pipeline/{vacuum_,}advance_e_pipeline.{h,cc}
       + remote.c:61-134 + local.c:50-122,224-264 */
void orc_advance_e( vpb_field_array_t * fa, float frac );
/* F12 src/field_advance/standard/pipeline/{vacuum_,}energy_f_pipeline.{h,c} */
void orc_energy_f( double * en /*[6]*/, const vpb_field_array_t * fa );

/* N1: divergence cleaning, charge synchronisation, shared-face synchronisation
   (sfa_private.h:256-373; local.c; remote.c:137-415,534-624; rho_p.cc:22-112)
   and N2's field initialisation kernels. */
/* 8(f) N3: hydro moments for the dump path */
void   orc_clear_hydro_array( vpb_hydro_array_t * ha );                           /* hydro_array.c:52-56 */
void   orc_accumulate_hydro_p( vpb_hydro_array_t * ha, const vpb_species_t * sp,
                               const vpb_interpolator_array_t * ia );             /* hydro_p.c:26-167 */
void   orc_synchronize_hydro_array( vpb_hydro_array_t * ha );                     /* hydro_array.c:71-222 */
void   orc_clear_rhof( vpb_field_array_t * fa );                                  /* sfa.c:239-245 */
void   orc_accumulate_rho_p( vpb_field_array_t * fa, const vpb_species_t * sp );   /* rho_p.cc:22-112 */
void   orc_synchronize_rho( vpb_field_array_t * fa );                             /* remote.c:534-624 */
void   orc_compute_rhob( vpb_field_array_t * fa );                                /* {vacuum_,}compute_rhob_pipeline */
void   orc_compute_curl_b( vpb_field_array_t * fa );                              /* {vacuum_,}compute_curl_b_pipeline */
double orc_synchronize_tang_e_norm_b( vpb_field_array_t * fa );                   /* remote.c:298-415 */
void   orc_compute_div_e_err( vpb_field_array_t * fa );                           /* {vacuum_,}compute_div_e_err_pipeline */
double orc_compute_rms_div_e_err( const vpb_field_array_t * fa );                 /* compute_rms_div_e_err_pipeline.c */
void   orc_clean_div_e( vpb_field_array_t * fa );                                 /* {vacuum_,}clean_div_e_pipeline */
void   orc_compute_div_b_err( vpb_field_array_t * fa );                           /* compute_div_b_err_pipeline.cc */
double orc_compute_rms_div_b_err( const vpb_field_array_t * fa );                 /* compute_rms_div_b_err_pipeline.c */
void   orc_clean_div_b( vpb_field_array_t * fa );                                 /* clean_div_b_pipeline.cc */
/* allreduce hook for the rms / mismatch diagnostics of multi-rank tests */
void   orc_set_allsum( void (*fn)( double * v, int n ) );

/* F14 src/boundary/boundary_p.cc:36-518 (absorb/reflect/periodic/remote; no
       custom handlers) and F15 src/species_advance/standard/rho_p.cc:125-171 */
void orc_boundary_p( vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa );
/* ... with the user handler list (boundary_p.cc:81-89, 236-253, 426-432) */
void orc_boundary_p_pbc( vpb_particle_bc_t * pbc_list, vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa );
/* a deterministic handler for tests (params: int[6] hit counters or NULL) */
int  orc_fixture_bounce_handler( void * params, vpb_species_t * sp, vpb_particle_t * p, vpb_particle_mover_t * pm,
                                 vpb_particle_injector_t * pi, int max_pi, int face );
/* src/collision/pipeline/langevin_pipeline.c:20-42, 99-108: u = decay u + drive n for every momentum
   component; normals = 3 deviates per particle in particle order (what frandn would have returned),
   or NULL for zeros (the deterministic part).  Returns nothing when step % interval != 0, like the reference. */
void orc_apply_langevin( vpb_langevin_t * l, const float * normals );
void orc_accumulate_rhob( vpb_field_t * f, const vpb_particle_t * p, const vpb_grid_t * g, float qsp );

/* Material coefficient table, src/field_advance/standard/sfa.c:55-151.
   Fills mc[0..n) for the list (ids 0..n-1); returns n. */
int  orc_material_coefficients( vpb_material_coefficient_t * mc, int max_mc,
                                const vpb_material_t * m_list, const vpb_grid_t * g, float damp );

#ifdef __cplusplus
}
#endif
#endif
