// vpb_runtime.h -- device runtime underneath the C-ABI entry points.
//
// One process drives one B200.  The host driver (the reference's
// vpic_simulation::advance, src/vpic/advance.cc:15-208, or our own
// vpb_sim_* loop) owns plain host structs; this runtime keeps a device MIRROR
// for each of them, keyed by the host struct's address, and the entry points
// launch kernels on the mirrors.  There is no CPU fallback anywhere: if CUDA is
// unusable every entry point dies with the reference's ERROR convention
// (src/util/util_base.h:267-273: "Error at file(line)[rank]" then exit(1)).
#pragma once

#include <cuda_runtime.h>
#include <cstdio>
#include <ctime>
#include <cstdlib>
#include <cstdint>
#include <unordered_map>
#include <vector>
#include "vpic_b200/abi.h"
#include "vpic_b200/vpic_b200.h"

namespace vpb {

extern int g_world_rank, g_world_size;
extern bool g_world_explicit;

[[noreturn]] void fatal( const char * file, int line, const char * fmt, ... );
void warning( const char * file, int line, const char * fmt, ... );

#define VPB_ERROR(...)   ::vpb::fatal( __FILE__, __LINE__, __VA_ARGS__ )
#define VPB_WARNING(...) ::vpb::warning( __FILE__, __LINE__, __VA_ARGS__ )
#define VPB_CUDA(call) do { cudaError_t e_ = (call); if( e_!=cudaSuccess ) \
  ::vpb::fatal( __FILE__, __LINE__, "CUDA error %s: %s", cudaGetErrorName(e_), cudaGetErrorString(e_) ); } while(0)
#define VPB_CHECK_LAUNCH() VPB_CUDA( cudaGetLastError() )

// ---- interior layouts on the device ---------------------------------------

// Fields are kept as structure-of-arrays: 16 float planes of nvp voxels each
// (nvp = nv rounded up to 32 so every plane starts 128-B aligned).  The order
// of the planes is the order of the floats in field_t.
enum FieldComp { F_EX=0, F_EY, F_EZ, F_DIVE, F_CBX, F_CBY, F_CBZ, F_DIVB,
                 F_TCAX, F_TCAY, F_TCAZ, F_RHOB, F_JFX, F_JFY, F_JFZ, F_RHOF, F_NCOMP };

// Interpolator: the 18 coefficients of interpolator_t, in the reference's
// order, padded to IP_STRIDE floats per voxel.  96 B = three 32 B sectors, so
// one voxel's record is fetched with three 256-bit loads (LDG.256, new on
// sm_100) and never straddles a sector; the host layout is 80 B
// (interpolator_t) and is converted at the coherence points only.
constexpr int IP_STRIDE = 24;   // floats per voxel on the device
constexpr int IP_HOST   = 20;   // floats per voxel in the host struct (80 B)
constexpr int ACC_STRIDE = 12;  // floats per voxel (48 B, reference layout)

// Particle boundary action per box face, derived from g->neighbor.
enum FaceAction : int { FACE_WRAP = 0, FACE_REFLECT = 1, FACE_EXIT = 2 };

struct GridDev {             // passed by value to kernels
  int nx, ny, nz;
  int sx, sy, sz, nv, nvp;
  int face[6];               // FaceAction for -x,-y,-z,+x,+y,+z
  const int32_t * neighbor;  // optional explicit table (6*nv), else nullptr
  // exact division of a voxel index (< 2^31) by sy, ny+2 and sz without the
  // ~20-instruction integer divide: q = umulhi( v, m ) >> sh  (see fast_div)
  unsigned m_sy, sh_sy, m_ny2, sh_ny2, m_sz, sh_sz;
};

// Magic pair for d >= 2: l = ceil(log2 d), m = ceil( 2^(31+l) / d ), sh = l-1.
// m*d - 2^(31+l) < d <= 2^l, so for every v < 2^31 the product error stays
// below 2^(31+l) and ( v*m ) >> (31+l) is floor( v/d ) exactly.
inline void fast_div_magic( unsigned d, unsigned & m, unsigned & sh ) {
  unsigned l = 0; while( ( 1ull<<l )<d ) l++;
  m = (unsigned)( ( ( 1ull<<( 31+l ) ) + d - 1 )/d ); sh = l-1;
}
#ifdef __CUDACC__
__device__ __forceinline__ int fast_div( int v, unsigned m, unsigned sh ) { return (int)( __umulhi( (unsigned)v, m ) >> sh ); }
#endif

// Encoding of the explicit neighbour table: >=0 local voxel, NB_REFLECT, and
// for everything that ends the local move (any other negative value): the
// host table's own code -- NB_ABSORB, a user handler id (<= -3,
// boundary_p.cc:236-253) -- or NB_REMOTE / NB_UNKNOWN.  The arithmetic path
// only says NB_EXIT and the face's kind decides (GridMirror::pkind).
constexpr int32_t NB_REFLECT = -1;
constexpr int32_t NB_ABSORB  = -2;
constexpr int32_t NB_EXIT    = -2;
constexpr int32_t NB_REMOTE  = INT32_MIN;
constexpr int32_t NB_UNKNOWN = INT32_MIN+1;

// ---- mirrors ---------------------------------------------------------------

// What boundary_p does with a particle that left through a box face
// (src/boundary/boundary_p.cc:196-260).
enum ParticleFaceKind : int { PK_LOCAL = 0,    // wrap / reflect: move_p never hands these over
                              PK_ABSORB = 1,   // absorb_particles: accumulate_rhob, drop
                              PK_REMOTE = 2,   // neighbouring rank: send an injector
                              PK_CUSTOM = 3 }; // user handler: a host callback of the host program (boundary_private.h:14-33)

struct GridMirror {
  const vpb_grid_t * host = nullptr;
  GridDev dev{};
  int32_t * d_neighbor = nullptr;
  int fbc[6];                // field boundary per face: rank (>=0) or local code (<0)
  int pkind[6];              // ParticleFaceKind per face
  int nb_n[6];               // for PK_REMOTE: the neighbour brick's cell count along the face axis
  long long pcode[6];        // for PK_CUSTOM: the face's neighbor[] code (user handler id, <= -3)
  bool per_voxel = false;    // set_region_bc etc.: boundary codes anywhere in the box, read from d_neighbor
  bool any_absorb = false, any_custom = false;   // per_voxel: which codes the table holds
  bool neighbors_final = false;                  // rebuilt once more at the first particle call (see particle_grid)
};

struct FieldMirror {
  vpb_field_array_t * host = nullptr;
  GridMirror * grid = nullptr;
  float   * d = nullptr;           // 16 * nvp floats
  int16_t * d_mat = nullptr;       // 8 * nvp material ids (only if n_mc>1)
  vpb_material_coefficient_t * d_mc = nullptr;
  int   n_mc = 0;
  float * d_stage = nullptr;       // AoS staging for host<->device conversion
  bool stage_has_ids = false;      // d_stage holds the host's material-id bytes (kept across downloads)
  double * d_en = nullptr;         // 6 doubles + scratch for energy_f
  float * d_face = nullptr;        // packed face buffers for exchanges
  size_t face_floats = 0;
  bool dev_valid = false;          // device copy holds the latest data
  bool host_valid = true;          // host copy holds the latest data
  float * comp( int c ) const { return d + (size_t)c*grid->dev.nvp; }
};

struct InterpMirror {
  vpb_interpolator_array_t * host = nullptr;
  GridMirror * grid = nullptr;
  float * d = nullptr;             // nv * IP_STRIDE floats
  float * d_stage = nullptr;       // nv * IP_HOST floats, host layout, for host<->device conversion
  bool dev_valid = false, host_valid = true;
};

struct AccMirror {
  vpb_accumulator_array_t * host = nullptr;
  GridMirror * grid = nullptr;
  float * d = nullptr;             // nv * 12 floats == host copy 0
  float * d_priv = nullptr;        // PARITY mode only: copies 1..n_priv (the reference's per-pipeline accumulators)
  int n_priv = 0;
  bool dev_valid = false, host_valid = true;
};

struct HydroMirror {
  vpb_hydro_array_t * host = nullptr;
  GridMirror * grid = nullptr;
  float * d = nullptr;             // nv * 16 floats == host hydro_t layout
  float * d_face = nullptr;        // [send_lo | send_hi | recv_lo | recv_hi] for synchronize_hydro_array
  size_t face_floats = 0;
  bool dev_valid = false, host_valid = true;
};

struct SpeciesMirror {
  vpb_species_t * host = nullptr;
  GridMirror * grid = nullptr;
  vpb_particle_t * d_p = nullptr;  size_t cap_p = 0;
  vpb_particle_mover_t * d_pm = nullptr; size_t cap_pm = 0;
  int * d_partition = nullptr;     // nv+1 ints
  int * d_partition_alt = nullptr; // the segment sort writes the new partition here while it reads the old one, then swaps
  bool home_valid = false;         // the array has the order of the last sort_p: position i is in the home segment d_partition gives it
  int home_np = 0;                 // particles at that sort (later ones were appended by boundary_p)
  int * d_counters = nullptr;      // [0]=nm  [1]=n_ignored  [2..] scratch
  int np = 0, nm = 0;              // device-side truth while resident
  bool movers_sorted = true;       // d_pm ascending in particle index
  bool dev_valid = false, host_valid = true;
  bool partition_dev_valid = false;
  const vpb_particle_t * registered_p = nullptr; size_t registered_bytes = 0;
  // Lazy sort_p (RESIDENT coherence + FAST deposit, vpb_sort.cu): the sort has been carried out on
  // indices only -- slot d of the sorted order is d_p[ d_perm[d] ] -- and the next advance_p gathers
  // through d_perm while it pushes.  Anything else that looks at the particles first makes the order
  // real (materialize_sort, called from to_device / to_host).
  int * d_perm = nullptr; size_t cap_perm = 0; bool perm_pending = false;
  // Voxel index of every particle as the last advance_p left it (written by the launch that precedes
  // a sort_p); valid until anything else moves, adds or removes particles (mark_dev_written).
  int * d_keys = nullptr; size_t cap_keys = 0; bool keys_valid = false;
  bool keys_patchable = false;       // d_keys is right for every particle that is not on the mover list (boundary_p patches the rest)
};

struct Runtime {
  bool initialised = false;
  int device = 0;
  cudaStream_t stream = nullptr;      // compute
  cudaStream_t copy_stream[2] = { nullptr, nullptr };
  int coherence = VPB_COHERENCE_STRICT;
  int deposit_mode = VPB_DEPOSIT_FAST;
  int ref_pipelines = 1;              // PARITY: the reference's N_PIPELINE (--tpp) whose summation order is reproduced
  int sm_count = 148;
  // scratch shared by all species (sort destination, keys, scan space)
  void * d_scratch = nullptr; size_t scratch_bytes = 0;
  void * h_pinned = nullptr;  size_t pinned_bytes = 0;
  long long launches = 0;             // kernels launched (for bench's gpu_launches)
  long long seg_sorts = 0;            // of which: segment sorts (deterministic ranks from the last sort's partition)
  long long sorts_from_keys = 0;      // lazy sort_p calls that took the voxel keys advance_p / boundary_p left (4 B per particle read)
  long long h2d_bytes = 0, d2h_bytes = 0;   // host<->device traffic issued by the entry points
  std::unordered_map<const void *, size_t>          pinned_host;   // host arrays registered with cudaHostRegister
  std::unordered_map<const void *, int>             host_touches;  // RESIDENT: transfers seen per host array (pin_host)
  std::unordered_map<const void *, GridMirror *>    grids;
  std::unordered_map<const void *, FieldMirror *>   fields;
  std::unordered_map<const void *, InterpMirror *>  interps;
  std::unordered_map<const void *, AccMirror *>     accs;
  std::unordered_map<const void *, SpeciesMirror *> species;
  std::unordered_map<const void *, HydroMirror *>   hydros;
};

Runtime & rt();
void ensure_init();
void pin_host( const void * ptr, size_t bytes );   // page-lock a host array the host program owns (idempotent, best effort)
void unpin_host( const void * ptr );
void unpin_all();
// write tracking of host arrays (vpb_track.cu; coherence mode TRACKED = STRICT + skipping uploads of arrays the host did not write)
bool track_enabled();
void track_enable( bool on );
void track_arm( const void * p, size_t bytes );      // host array == device mirror right now: write-protect it
bool track_clean( const void * p, size_t bytes );    // not written by the host program since it was armed
void track_disarm( const void * p );                 // give the pages back (before freeing / resizing, or forgetting the object)
void track_forget_all();
void * scratch( size_t bytes );        // grow-only device scratch
void * pinned( size_t bytes );         // grow-only pinned host scratch

GridMirror    * mirror( const vpb_grid_t * g );
void build_neighbors( GridMirror * m );          // (re)derive everything that comes from g->neighbor
// The host program may edit g->neighbor between creating its arrays and the
// first step (deck/wrapper.h set_region_bc); the particle kernels therefore
// re-read the table at their first call.
inline GridMirror * particle_grid( GridMirror * m ) { if( !m->neighbors_final ) { build_neighbors( m ); m->neighbors_final = true; } return m; }
FieldMirror   * mirror( vpb_field_array_t * fa );
InterpMirror  * mirror( vpb_interpolator_array_t * ia );
AccMirror     * mirror( vpb_accumulator_array_t * aa );
SpeciesMirror * mirror( vpb_species_t * sp );
HydroMirror   * mirror( vpb_hydro_array_t * ha );
void release( HydroMirror * m );

// Coherence helpers.  to_device(): make the device copy current (uploads if the
// host copy is the valid one, or always in STRICT mode).  to_host(): the
// reverse.  mark_dev_written(): the device copy was modified by a kernel.
void to_device( FieldMirror * m );   void to_host( FieldMirror * m );
void to_device( InterpMirror * m );  void to_host( InterpMirror * m );
void to_device( AccMirror * m );     void to_host( AccMirror * m );
void to_device( SpeciesMirror * m, bool keep_lazy_sort = false ); void to_host( SpeciesMirror * m );
void materialize_sort( SpeciesMirror * m );            // carry out a pending lazy sort_p (no-op otherwise)
vpb_particle_t * sort_spare( size_t min_particles );   // the spare particle buffer shared by all species
void sort_swap_spare( SpeciesMirror * m );             // m->d_p <-> the spare buffer (after an out-of-place pass)
void release_lazy_sort( SpeciesMirror * m );
void to_device( HydroMirror * m );   void to_host( HydroMirror * m );
void movers_to_host( SpeciesMirror * m );
template<class M> inline void mark_dev_written( M * m ) { m->dev_valid = true; m->host_valid = false; }
inline void mark_dev_written( SpeciesMirror * m ) { m->dev_valid = true; m->host_valid = false; m->keys_valid = m->keys_patchable = false; m->home_valid = false; }
// In STRICT mode every entry point finishes by pushing what it wrote back.
template<class M> inline void finish_write( M * m ) {
  mark_dev_written( m );
  if( rt().coherence==VPB_COHERENCE_STRICT ) to_host( m );
}

// cross-file internals
struct PushCoef;
void advance_p_parity( SpeciesMirror *, AccMirror *, InterpMirror *, const PushCoef & );
void reduce_accumulators_parity( AccMirror * );
void accumulate_rho_p_parity( FieldMirror * mf, SpeciesMirror * ms, float q_8V );   // reference summation order
struct HydroCoef;
void accumulate_hydro_p_parity( HydroMirror * mh, SpeciesMirror * ms, InterpMirror * mi, const HydroCoef & k );   // reference summation order
void energy_f_parity( double * en6, FieldMirror * m );           // local (pre-allreduce) energies, reference summation order
double energy_p_parity( SpeciesMirror * ms, InterpMirror * mi, float qdt_2mc, float msp );
// counting-sort front end shared by sort_p and the PARITY deposit: start[0..nbin] = first slot of each bin,
// perm[0..n) = element indices grouped by bin (order inside a bin arbitrary)
void bin_indices( const int * keys, int n, int nbin, int * start, int * perm );
void release_parity_buffers();
void sort_movers_by_index( SpeciesMirror * );
void comm_exchange_faces( FieldMirror * m, int X, float * slo, float * shi, float * rlo, float * rhi, int nfloat );
void comm_sendrecv( int rank_lo, int rank_hi, const void * slo, size_t nslo, const void * shi, size_t nshi,
                    void * rlo, size_t nrlo, void * rhi, size_t nrhi );   // device buffers, byte counts; rank<0 = no neighbour
// Mailbox slots of one fused exchange (vpb_comm.cu: comm_link_reserve); index 0 = low side, 1 = high side.
struct CommLink {
  char * dst[2] = { nullptr, nullptr };  unsigned * pub_flag[2] = { nullptr, nullptr };        unsigned pub_seq[2] = { 0, 0 };
  const char * src[2] = { nullptr, nullptr }; const unsigned * wait_flag[2] = { nullptr, nullptr }; unsigned wait_seq[2] = { 0, 0 };
  unsigned * done = nullptr;             // zeroed block counters (64) for "last block publishes"
};
bool comm_link_reserve( int rank_lo, int rank_hi, size_t nbytes, CommLink * out );
void comm_allsum_d( double * v, int n );
bool comm_ready();
void comm_shutdown();
void release_boundary_buffers();
void host_resize_particles( vpb_species_t * sp, size_t n, size_t keep );   // boundary_p.cc:346-394 for host-resident arrays; keep = elements whose contents matter
void host_resize_movers( vpb_species_t * sp, size_t n, size_t keep );
void exclusive_scan( const int * in, int * out, int * out2, int * tmp, int n );
// field boundary machinery shared by vpb_fields.cu and vpb_divclean.cu
bool face_is_remote( const GridMirror * gm, int X, int side );
int  face_bc( const GridMirror * gm, int X, int side );
void exchange_axis( FieldMirror * m, int X, int nfloat, float ** from_lo, float ** from_hi, float * base = nullptr );   // packed faces in base (default m->d_face)
void fill_ghost_tang_b( FieldMirror * m );
void local_adjust_norm_b( FieldMirror * m );
void local_adjust_tang_e( FieldMirror * m );
void release_sort_buffers();
__global__ void k_sum_partials( double * out, const double * partial, int nblock, int ncomp, double scale );

// every host<->device copy of the library goes through these two (byte accounting for bench.py's e2e)
// VPIC_B200_TRACE_TRANSFERS=1: one line on stderr per transfer of 1 MB or more (which arrays a coherence mode moves)
inline bool trace_transfers() { static const bool on = getenv( "VPIC_B200_TRACE_TRANSFERS" )!=nullptr; return on; }
inline void h2d( void * dst, const void * src, size_t bytes, cudaStream_t st ) {
  VPB_CUDA( cudaMemcpyAsync( dst, src, bytes, cudaMemcpyHostToDevice, st ) ); rt().h2d_bytes += (long long)bytes;
  if( bytes>=( 1u<<20 ) && trace_transfers() ) fprintf( stderr, "[vpic_b200] h2d %zu bytes from %p\n", bytes, src );
}
inline void d2h( void * dst, const void * src, size_t bytes, cudaStream_t st ) {
  VPB_CUDA( cudaMemcpyAsync( dst, src, bytes, cudaMemcpyDeviceToHost, st ) ); rt().d2h_bytes += (long long)bytes;
  if( bytes>=( 1u<<20 ) && trace_transfers() ) fprintf( stderr, "[vpic_b200] d2h %zu bytes to %p\n", bytes, dst );
}

inline int div_up( long long a, long long b ) { return (int)( (a+b-1)/b ); }

// VPIC_B200_TRACE_CALLS=ms: one line on stderr for every traced entry point call that took at least that many
// milliseconds of host wall time (first-touch uploads, page-locking, allocations: what a deck's start-up pays)
struct CallTrace {
  const char * name; timespec t0; double min_ms;
  explicit CallTrace( const char * n ) : name( n ) {
    static const double thr = [](){ const char * e = getenv( "VPIC_B200_TRACE_CALLS" ); return e ? atof( e ) : -1.0; }();
    min_ms = thr; if( min_ms>=0 ) clock_gettime( CLOCK_MONOTONIC, &t0 );
  }
  ~CallTrace() {
    if( min_ms<0 ) return;
    timespec t1; clock_gettime( CLOCK_MONOTONIC, &t1 );
    const double ms = ( t1.tv_sec-t0.tv_sec )*1e3 + ( t1.tv_nsec-t0.tv_nsec )*1e-6;
    if( ms>=min_ms ) fprintf( stderr, "[vpic_b200] %s: %.2f ms\n", name, ms );
  }
};
#define VPB_TRACE_CALL( name ) vpb::CallTrace vpb_call_trace_( name )

} // namespace vpb
