/* vpic_b200.h -- C ABI of libvpic_b200.so, the B200 (sm_100a) implementation
 * of VPIC's per-step PIC hot path.
 *
 * Part 1 exports THE REFERENCE'S OWN SYMBOLS for the path (same name, same
 * argument meaning, same error behaviour), so the library links underneath the
 * unmodified vpic_simulation driver and input decks in place of the
 * reference's species_advance / sf_interface / field_advance translation
 * units.  Struct types are our byte-compatible restatements
 * from abi.h; in C the type names are irrelevant to the ABI.
 *
 * Part 2 is the small control surface a host needs on top: coherence between
 * the host structs and the device mirrors, deposit mode, multi-GPU comm.
 *
 * Part 3 is a stand-alone host layer (grid / species / array constructors and
 * a time step) mirroring the reference's constructors, for programs that do
 * not link the reference at all (bench.py, tests).
 *
 * All entry points must be called from one host thread (the reference's rule:
 * src/util/pipelines/pipelines_thread.c:529).  Failures follow the reference:
 * print "Error at file(line)[rank]: ..." and exit(1); soft failures warn.
 */
#ifndef VPIC_B200_H
#define VPIC_B200_H

#include "vpic_b200/abi.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ======================= Part 1: reference symbols ======================== */

/* src/species_advance/species_advance.h:65-66 */
void sort_p( vpb_species_t * sp );
/* src/species_advance/species_advance.h:73-76 */
void advance_p( vpb_species_t * sp, vpb_accumulator_array_t * aa, const vpb_interpolator_array_t * ia );
/* src/species_advance/species_advance.h:118-120 */
double energy_p( const vpb_species_t * sp, const vpb_interpolator_array_t * ia );
/* src/species_advance/species_advance.h:88-104 */
void center_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia );
void uncenter_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia );
/* src/sf_interface/sf_interface.h:99-101 */
void load_interpolator_array( vpb_interpolator_array_t * ia, const vpb_field_array_t * fa );
/* src/sf_interface/sf_interface.h:147-148 */
void clear_accumulator_array( vpb_accumulator_array_t * aa );
/* src/sf_interface/sf_interface.h:158-159 */
void reduce_accumulator_array( vpb_accumulator_array_t * aa );
/* src/sf_interface/sf_interface.h:172-174 */
void unload_accumulator_array( vpb_field_array_t * fa, const vpb_accumulator_array_t * aa );
/* src/boundary/boundary.h:35-39 (pbc_list: opaque particle_bc_t list, may be NULL) */
void boundary_p( void * pbc_list, vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa );

/* src/field_advance/field_advance.h:233-236.  Creates a field array whose
   kernel table (field_advance.h:170-218) points at the functions below. */
vpb_field_array_t * new_standard_field_array( vpb_grid_t * g, const vpb_material_t * m_list, float damp );
void delete_field_array( vpb_field_array_t * fa );

/* The kernel-table entries themselves (global symbols: the reference
   checkpoints function pointers by name, src/field_advance/field_advance.c:17-58). */
void vpic_b200_delete_fa( vpb_field_array_t * fa );
/* checkpoint / restore hooks registered with the reference's checkpt service when the
   host program has one (sfa.c:161-188; src/util/checkpt/checkpt.h:150-175) */
void   vpic_b200_checkpt_fa( const void * fa );
void * vpic_b200_restore_fa( void );
void vpic_b200_advance_b( vpb_field_array_t * fa, float frac );       /* advance_b.cc:10-19     */
void vpic_b200_advance_e( vpb_field_array_t * fa, float frac );       /* advance_e.cc:10-25     */
void vpic_b200_energy_f( double * en, const vpb_field_array_t * fa ); /* energy_f.c:12-23       */
void vpic_b200_clear_jf( vpb_field_array_t * fa );                    /* sfa.c:231-237          */
void vpic_b200_synchronize_jf( vpb_field_array_t * fa );              /* remote.c:417-508       */
void vpic_b200_clear_rhof( vpb_field_array_t * fa );                  /* sfa.c:239-245          */
void vpic_b200_synchronize_rho( vpb_field_array_t * fa );             /* remote.c:534-624       */
void vpic_b200_compute_rhob( vpb_field_array_t * fa );                /* {vacuum_,}compute_rhob.c:10-21 */
void vpic_b200_compute_curl_b( vpb_field_array_t * fa );              /* {vacuum_,}compute_curl_b.cc:10-19 */
double vpic_b200_synchronize_tang_e_norm_b( vpb_field_array_t * fa ); /* remote.c:298-415       */
void vpic_b200_compute_div_e_err( vpb_field_array_t * fa );           /* {vacuum_,}compute_div_e_err.c:10-21 */
double vpic_b200_compute_rms_div_e_err( const vpb_field_array_t * fa ); /* compute_rms_div_e_err.c:12-23 */
void vpic_b200_clean_div_e( vpb_field_array_t * fa );                 /* {vacuum_,}clean_div_e.c:10-19 */
void vpic_b200_compute_div_b_err( vpb_field_array_t * fa );           /* compute_div_b_err.cc:10-19 */
double vpic_b200_compute_rms_div_b_err( const vpb_field_array_t * fa ); /* compute_rms_div_b_err.c:12-23 */
void vpic_b200_clean_div_b( vpb_field_array_t * fa );                 /* clean_div_b.cc:10-19   */

/* Charge density of a species onto rhof (reference symbol, called by the
   driver between clear_rhof and synchronize_rho, src/vpic/advance.cc:144-146;
   src/species_advance/standard/rho_p.cc:22-112). */
void accumulate_rho_p( vpb_field_array_t * fa, const vpb_species_t * sp );

/* Langevin thermostat: src/collision/langevin.h:14-15, pipeline/langevin_pipeline.c:44-112.
   u = decay u + drive N(0,1) per momentum component, decay = exp(-nu dt interval),
   drive = sqrt( (1 - exp(-2 nu dt interval)) kT / (m c) ); applied when step % interval == 0.  The reference's
   apply_langevin (langevin.c:12-24, the function its collision_op_t holds) calls this symbol, so a program that
   links the library ahead of the reference runs its langevin() operators on the device.  The normal deviates
   come from a counter-based generator keyed by (seed, rank, species id, step, particle index) instead of the
   host program's rng_pool: same distribution, different stream -- statistical parity, not bit parity. */
void apply_langevin_pipeline( vpb_langevin_t * l );
void vpic_b200_set_collision_seed( uint64_t seed );   /* default 0x5eed */

/* Hydro moments for the dump path.
   Replace src/species_advance/species_advance.h accumulate_hydro_p
   (standard/hydro_p.c:26-167) and src/sf_interface/sf_interface.h:214-225
   clear_hydro_array / synchronize_hydro_array (hydro_array.c:52-222). */
void accumulate_hydro_p( vpb_hydro_array_t * ha, const vpb_species_t * sp, const vpb_interpolator_array_t * ia );
void clear_hydro_array( vpb_hydro_array_t * ha );
void synchronize_hydro_array( vpb_hydro_array_t * ha );
/* standalone construction; a reference build keeps its own new_/delete_hydro_array */
vpb_hydro_array_t * vpic_b200_new_hydro_array( vpb_grid_t * g );
void vpic_b200_delete_hydro_array( vpb_hydro_array_t * ha );

/* ======================= Part 2: control surface ========================== */

enum { VPB_COHERENCE_STRICT = 0,   /* every call: host structs in, host structs out  */
       VPB_COHERENCE_RESIDENT = 1, /* state lives on the device between calls; the
                                      host syncs explicitly with the calls below     */
       VPB_COHERENCE_TRACKED = 2   /* STRICT, minus the uploads of arrays the host program has not
                                      written since the library last saw them: after every upload /
                                      download the array's pages are write-protected and a store
                                      into them marks the array dirty (SIGSEGV handler, mprotect).
                                      The host copy stays current after every call exactly as in
                                      STRICT mode.  Not for programs that fill their arrays with
                                      read()/fread(): system calls fail with EFAULT on protected
                                      pages instead of faulting (vpb_track.cu).                  */ };
enum { VPB_DEPOSIT_FAST = 0,       /* warp-reduced + atomics, FMA allowed            */
       VPB_DEPOSIT_PARITY = 1      /* reference summation order, bit-exact currents  */ };

int  vpic_b200_init( int device, int world_rank, int world_size ); /* returns SM count */
/* rank / size only (host state for the grid partitioners; no CUDA needed) */
void vpic_b200_set_world( int world_rank, int world_size );
/* Multi-GPU transport (replaces src/util/mp for device buffers): one NCCL
   communicator, one process per GPU.  Rank 0 obtains the 128-byte id, the host
   program distributes it (MPI_Bcast in a reference build, torch.distributed in
   the tests), every rank calls comm_init after vpic_b200_init. */
int  vpic_b200_comm_unique_id( void * out, int nbytes );           /* returns bytes written */
void vpic_b200_comm_init( const void * id_bytes, int nbytes );
int  vpic_b200_comm_ready( void );
/* The same transport WITHOUT NCCL (and therefore also with several ranks on one GPU): every rank
   obtains its 80-byte mailbox record (cudaIpc handle), the host program all-gathers the records
   (rank order), every rank installs the set.  Exchanges, migration and the all-reduce (summed in
   rank order, so every rank gets the same bits) then run on our put kernels over peer memory.  A
   host program built from the reference needs neither call: the first exchange bootstraps itself
   over the program's own mp_allgather_i (src/util/mp/mp.h:76-79). */
int  vpic_b200_comm_mailbox_record( void * out, int nbytes );      /* returns bytes written (80) */
void vpic_b200_comm_init_handles( const void * all_records, int nbytes_each );
/* 1 = face exchanges / particle migration use the put transport (kernels storing
   into the neighbours' IPC-mapped mailboxes over NVLink), 0 = ncclSend/ncclRecv
   (VPIC_B200_TRANSPORT=nccl, or the mailbox could not be exported). */
int  vpic_b200_comm_transport( void );
/* Diagnostic: right after sort_p, the number of particles that are out of voxel
   order or outside the segment sp->partition gives their voxel (0 expected; -1 if
   the device holds no partition).  Lets tests check sort_p at sizes no CPU oracle reaches. */
long long vpic_b200_sort_violations( vpb_species_t * sp );
/* Diagnostic: how many sort_p calls of this process read the 4-byte voxel keys the
   preceding advance_p wrote (and boundary_p patched) instead of the particle records. */
long long vpic_b200_sorts_from_keys( void );
/* ... and how many of those were segment sorts (deterministic ranks from the last sort's
   partition; VPIC_B200_SEG_SORT=0 disables them). */
long long vpic_b200_segment_sorts( void );
void vpic_b200_finalize( void );
/* frees the grow-on-demand device caches (sort_p spare buffer, scratch, boundary_p
   message buffers); objects and their mirrors stay valid */
void vpic_b200_trim( void );
void vpic_b200_set_coherence( int mode );
void vpic_b200_set_deposit_mode( int mode );
/* PARITY only: the reference's pipeline count (--tpp) whose summation order the
   energy diagnostics reproduce (advance_p takes it from aa->n_pipeline). Default 1. */
void vpic_b200_set_reference_pipelines( int n_pipeline );
int  vpic_b200_abi_check( int sizeof_interpolator, int sizeof_accumulator, int sizeof_field,
                          int sizeof_particle, int sizeof_grid, int sizeof_species ); /* 0 = ok */
/* obj = any species_t / field_array_t / interpolator_array_t / accumulator_array_t */
void vpic_b200_sync_to_host( void * obj );     /* device -> host struct arrays      */
void vpic_b200_sync_from_host( void * obj );   /* host struct arrays -> device       */
void vpic_b200_forget( void * obj );           /* drop the mirror (before free())    */
/* g->neighbor / g->bc were edited after the first particle step (the edits a
   deck makes before its first step -- set_region_bc, set_domain_particle_bc --
   are picked up without this call: the table is read at the first advance_p). */
void vpic_b200_grid_changed( const void * grid );
void vpic_b200_device_synchronize( void );
long long vpic_b200_launch_count( void );      /* kernels launched so far            */
void vpic_b200_transfer_bytes( long long * h2d_d2h ); /* bytes copied host->device, device->host so far */
void * vpic_b200_stream( void );               /* cudaStream_t the kernels run on    */
/* Device pointers of a mirror, for tests/benchmarks: which = 0 particles,
   1 movers, 2 partition (species); 0 SoA planes (field); 0 coefficients
   (interpolator); 0 accumulator. */
void * vpic_b200_device_ptr( void * obj, int which );

/* src/vpic/advance.cc:78-101: drop the particles whose movers are still unprocessed after
   num_comm_round boundary_p calls (voxel restored, charge accumulated to rhob, back-fill in
   reverse mover order), for host programs whose particles live on the device; a program built
   from the reference runs that loop itself on the host arrays.  Returns the particles removed. */
int vpic_b200_drop_unprocessed_movers( vpb_species_t * sp_list, vpb_field_array_t * fa );

/* Synthetic loader used by bench.py: fills sp with ppc particles per cell,
   uniform in each voxel, Maxwellian momenta (vth in units of c), weight w,
   from a counter-based generator, directly on the device (no host copy). */
void vpic_b200_load_uniform_maxwellian( vpb_species_t * sp, int ppc, float vth, float w, uint64_t seed );

/* NO COUNTERPART IN THE REFERENCE.
   energy_f (vacuum, one material) for callers that may not allocate (CUDA-graph
   capture, tight serving loops): the per-block partial sums live in a file-scope
   static device array instead of the cudaMalloc'ed scratch arena, so the call
   performs no cudaMalloc / cudaFree and touches no runtime state.
     d_en6       out, DEVICE pointer to 6 doubles (ex, ey, ez, cbx, cby, cbz energies, energy_f.c order)
     d_field_soa in, DEVICE pointer to the field mirror's SoA planes = vpic_b200_device_ptr( fa, 0 )
     nx, ny, nz  the grid's local cell counts (the planes' ghost layout follows from them)
     eps_rmu6    HOST pointer: epsx, epsy, epsz, rmux, rmuy, rmuz of the material
     scale       0.5 * eps0 * dV (what energy_f multiplies the sums with)
     stream      cudaStream_t the two kernels are launched on (NULL = default stream)
   Limit: nx*ny*nz <= 33 554 432 cells (384 KB of static device memory per GPU, held until
   exit); larger grids return -2.  Returns 0, or -1 bad args, -2 too large, -3 CUDA error
   (a message goes to stderr; the process is NOT terminated).  All calls share the one
   array: calls that overlap on the GPU race; the first call on a device must be made
   outside stream capture (it looks the array's device address up). */
int vpic_b200_energy_f_noalloc( double * d_en6, const float * d_field_soa, int nx, int ny, int nz,
                                const float * eps_rmu6, double scale, void * stream );

/* ======================= Part 3: stand-alone host layer =================== */

/* src/grid/grid_structors.c, partition.c:25-182, ops.c:18-211 */
vpb_grid_t * vpic_b200_new_grid( void );
void vpic_b200_delete_grid( vpb_grid_t * g );
void vpic_b200_partition_periodic_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                       double gx1, double gy1, double gz1,
                                       int gnx, int gny, int gnz, int gpx, int gpy, int gpz );
void vpic_b200_partition_absorbing_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                        double gx1, double gy1, double gz1,
                                        int gnx, int gny, int gnz, int gpx, int gpy, int gpz, int pbc );
void vpic_b200_partition_metal_box( vpb_grid_t * g, double gx0, double gy0, double gz0,
                                    double gx1, double gy1, double gz1,
                                    int gnx, int gny, int gnz, int gpx, int gpy, int gpz );
void vpic_b200_set_fbc( vpb_grid_t * g, int boundary, int fbc );
void vpic_b200_set_pbc( vpb_grid_t * g, int boundary, int pbc );
/* src/species_advance/species_advance.c:102-146; host_arrays=0 keeps p/pm
   device-only (sp->p == NULL), for problems larger than host memory. */
vpb_species_t * vpic_b200_new_species( const char * name, float q, float m, size_t max_np, size_t max_nm,
                                       int sort_interval, int sort_out_of_place, vpb_grid_t * g, int host_arrays );
void vpic_b200_delete_species( vpb_species_t * sp );
/* src/sf_interface/interpolator_array.cc:18-50, accumulator_array.c:52-73 */
vpb_interpolator_array_t * vpic_b200_new_interpolator_array( vpb_grid_t * g );
void vpic_b200_delete_interpolator_array( vpb_interpolator_array_t * ia );
vpb_accumulator_array_t * vpic_b200_new_accumulator_array( vpb_grid_t * g, int n_pipeline );
void vpic_b200_delete_accumulator_array( vpb_accumulator_array_t * aa );
/* one-material list helper: src/material/material.c */
vpb_material_t * vpic_b200_new_material( const char * name, float eps, float mu, float sigma );
vpb_material_t * vpic_b200_append_material( vpb_material_t * m, vpb_material_t ** m_list );   /* material.c:69-86 */

#ifdef __cplusplus
}
#endif
#endif /* VPIC_B200_H */
