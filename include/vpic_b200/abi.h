/* vpic_b200/abi.h -- binary layouts shared with the reference host code.
 *
 * The B200 library plugs in underneath the reference's unmodified driver and
 * input decks, so it has to read and write the very same plain-old-data
 * structures the host code owns.  This header is OUR restatement of those
 * layouts (no reference header is included); every struct cites the reference
 * definition it must stay byte-compatible with, and the static asserts at the
 * bottom pin size and field offsets.  tests/test_abi.py cross-checks them
 * against the reference's own sizeof/offsetof whenever the reference is
 * available.
 *
 * Layout variant: the reference changes interpolator_t / accumulator_t padding
 * when it is compiled with USE_V8_* or USE_V16_* (src/sf_interface/
 * sf_interface.h:27-53).  This library is built for the un-padded ("scalar" /
 * V4) layout: interpolator 80 B, accumulator 48 B.  vpic_b200_abi_check()
 * lets a host refuse to run against a differently padded libvpic.
 */
#ifndef VPIC_B200_ABI_H
#define VPIC_B200_ABI_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* src/util/pipelines/pipelines.h:10 */
enum { VPB_MAX_PIPELINE = 272 };

/* ---- particles: src/species_advance/species_advance_aos.h:21-52 ---------- */

typedef struct vpb_particle {
  float   dx, dy, dz;   /* voxel-relative position on [-1,1]                  */
  int32_t i;            /* local voxel index; 8*voxel+face while awaiting
                           boundary_p (species_advance_aos.h:23-29)           */
  float   ux, uy, uz;   /* normalised momentum gamma*beta                     */
  float   w;            /* weight (number of physical particles)              */
} vpb_particle_t;

typedef struct vpb_particle_mover {
  float   dispx, dispy, dispz; /* remaining displacement, half-cell units     */
  int32_t i;                   /* index of the particle in sp->p              */
} vpb_particle_mover_t;

typedef struct vpb_particle_injector {
  float   dx, dy, dz;  int32_t i;
  float   ux, uy, uz;  float   w;
  float   dispx, dispy, dispz;
  int32_t sp_id;
} vpb_particle_injector_t;

/* ---- grid: src/grid/grid.h:73-131 ---------------------------------------- */

enum {
  VPB_PEC_FIELDS        = -1,  /* grid.h:21-23 anti_symmetric/pec/metal       */
  VPB_SYMMETRIC_FIELDS  = -2,  /* grid.h:24                                   */
  VPB_PMC_FIELDS        = -3,  /* grid.h:25                                   */
  VPB_ABSORB_FIELDS     = -4,  /* grid.h:26                                   */
  VPB_REFLECT_PARTICLES = -1,  /* grid.h:29                                   */
  VPB_ABSORB_PARTICLES  = -2   /* grid.h:30                                   */
};

/* BOUNDARY(i,j,k), grid.h:16 */
#define VPB_BOUNDARY(i,j,k) (13+(i)+3*(j)+9*(k))
/* VOXEL(x,y,z,nx,ny,nz), grid.h:136 */
#define VPB_VOXEL(x,y,z,nx,ny,nz) ((x)+((nx)+2)*((y)+((ny)+2)*(z)))

typedef struct vpb_grid {
  float   dt, cvac, eps0;
  int64_t step;
  double  t0;
  float   x0, y0, z0;
  float   x1, y1, z1;
  int     nx, ny, nz;
  float   dx, dy, dz, dV;
  float   rdx, rdy, rdz, r8V;
  int     sx, sy, sz, nv;
  int     bc[27];
  int64_t *range;      /* (0:nproc) global voxel ranges                       */
  int64_t *neighbor;   /* (0:5, 0:nv-1) global neighbour ids or BC codes      */
  int64_t rangel, rangeh;
  void    *mp;         /* mp_t*, opaque here                                  */
} vpb_grid_t;

/* ---- species: src/species_advance/species_advance_aos.h:54-94 ------------ */

typedef struct vpb_species {
  char                 *name;
  float                 q, m;
  int                   np, max_np;
  vpb_particle_t       *p;
  int                   nm, max_nm;
  vpb_particle_mover_t *pm;
  int64_t               last_sorted;
  int                   sort_interval;
  int                   sort_out_of_place;
  int                  *partition;      /* nv+1 ints (species_advance.c:138)  */
  vpb_grid_t           *g;
  int32_t               id;
  struct vpb_species   *next;
} vpb_species_t;

/* ---- interpolator / accumulator: src/sf_interface/sf_interface.h:62-131 -- */

typedef struct vpb_interpolator {
  float ex, dexdy, dexdz, d2exdydz;
  float ey, deydz, deydx, d2eydzdx;
  float ez, dezdx, dezdy, d2ezdxdy;
  float cbx, dcbxdx;
  float cby, dcbydy;
  float cbz, dcbzdz;
  float _pad1[2];
} vpb_interpolator_t;

typedef struct vpb_interpolator_array {
  vpb_interpolator_t *i;
  vpb_grid_t         *g;
} vpb_interpolator_array_t;

typedef struct vpb_accumulator {
  float jx[4];   /* jx0@(0,-1,-1) jx1@(0,1,-1) jx2@(0,-1,1) jx3@(0,1,1)       */
  float jy[4];
  float jz[4];
} vpb_accumulator_t;

typedef struct vpb_accumulator_array {
  vpb_accumulator_t *a;
  int                n_pipeline;   /* copies 1..n_pipeline belong to threads  */
  int                stride;       /* POW2_CEIL(nv,2), accumulator_array.c:56 */
  vpb_grid_t        *g;
} vpb_accumulator_array_t;

/* ---- hydro moments: src/sf_interface/sf_interface.h:185-198 -------------- */

typedef struct vpb_hydro {
  float jx, jy, jz, rho;   /* current and charge density   */
  float px, py, pz, ke;    /* momentum and kinetic energy density */
  float txx, tyy, tzz;     /* stress diagonal              */
  float tyz, tzx, txy;     /* stress off-diagonal          */
  float _pad[2];           /* 64 B                         */
} vpb_hydro_t;

typedef struct vpb_hydro_array {
  vpb_hydro_t *h;
  vpb_grid_t  *g;
} vpb_hydro_array_t;

/* ---- Langevin collision operator: src/collision/langevin.h:6-13 ---------- */

typedef struct vpb_langevin {
  vpb_species_t * sp;
  void          * rp;        /* rng_pool_t * of the host program (not used on the device: counter-based generator) */
  float kT;
  float nu;
  int   interval;
} vpb_langevin_t;

/* ---- fields: src/field_advance/field_advance.h:152-229 ------------------- */

typedef int16_t vpb_material_id;

typedef struct vpb_field {
  float ex,   ey,   ez,   div_e_err;
  float cbx,  cby,  cbz,  div_b_err;
  float tcax, tcay, tcaz, rhob;
  float jfx,  jfy,  jfz,  rhof;
  vpb_material_id ematx, ematy, ematz, nmat;
  vpb_material_id fmatx, fmaty, fmatz, cmat;
} vpb_field_t;

struct vpb_field_array;

typedef struct vpb_field_advance_kernels {
  void   (*delete_fa)( struct vpb_field_array * fa );
  void   (*advance_b)( struct vpb_field_array * fa, float frac );
  void   (*advance_e)( struct vpb_field_array * fa, float frac );
  void   (*energy_f)( double * en, const struct vpb_field_array * fa );
  void   (*clear_jf)( struct vpb_field_array * fa );
  void   (*synchronize_jf)( struct vpb_field_array * fa );
  void   (*clear_rhof)( struct vpb_field_array * fa );
  void   (*synchronize_rho)( struct vpb_field_array * fa );
  void   (*compute_rhob)( struct vpb_field_array * fa );
  void   (*compute_curl_b)( struct vpb_field_array * fa );
  double (*synchronize_tang_e_norm_b)( struct vpb_field_array * fa );
  void   (*compute_div_e_err)( struct vpb_field_array * fa );
  double (*compute_rms_div_e_err)( const struct vpb_field_array * fa );
  void   (*clean_div_e)( struct vpb_field_array * fa );
  void   (*compute_div_b_err)( struct vpb_field_array * fa );
  double (*compute_rms_div_b_err)( const struct vpb_field_array * fa );
  void   (*clean_div_b)( struct vpb_field_array * fa );
} vpb_field_advance_kernels_t;

/* particle_bc_t, src/boundary/boundary_private.h:14-44: a user handler for
   particles that hit a face whose neighbor[] code is id (<= -3).  interact()
   is a HOST function of the host program; boundary_p calls it for every such
   particle in the reference's order and injects what it returns. */
struct vpb_species;
typedef int (*vpb_particle_bc_func_t)( void * params, struct vpb_species * sp, vpb_particle_t * p, vpb_particle_mover_t * pm,
                                       vpb_particle_injector_t * pi, int max_pi, int face );
typedef struct vpb_particle_bc {
  void * params;
  vpb_particle_bc_func_t interact;
  void (*delete_pbc)( struct vpb_particle_bc * pbc );
  int64_t id;
  struct vpb_particle_bc * next;
} vpb_particle_bc_t;

typedef struct vpb_field_array {
  vpb_field_t                *f;
  vpb_grid_t                 *g;
  void                       *params;     /* vpb_sfa_params_t* for ours       */
  vpb_field_advance_kernels_t kernel[1];
} vpb_field_array_t;

/* src/field_advance/standard/sfa_private.h:14-30 */
typedef struct vpb_material_coefficient {
  float decayx, drivex;
  float decayy, drivey;
  float decayz, drivez;
  float rmux, rmuy, rmuz;
  float nonconductive;
  float epsx, epsy, epsz;
  float pad[3];
} vpb_material_coefficient_t;

typedef struct vpb_sfa_params {
  vpb_material_coefficient_t *mc;
  int                         n_mc;
  float                       damp;
} vpb_sfa_params_t;

/* src/material/material.h:12-20 */
typedef struct vpb_material {
  char  *name;
  float  epsx, epsy, epsz;
  float  mux, muy, muz;
  float  sigmax, sigmay, sigmaz;
  float  zetax, zetay, zetaz;
  vpb_material_id id;
  struct vpb_material *next;
} vpb_material_t;

#ifdef __cplusplus
} /* extern "C" */
#endif

/* ---- layout pins --------------------------------------------------------- */
#if defined(__cplusplus)
#define VPB_SA(c,m) static_assert(c,m)
#else
#define VPB_SA(c,m) _Static_assert(c,m)
#endif

VPB_SA( sizeof(vpb_particle_t)==32,           "particle_t is 32 B" );
VPB_SA( sizeof(vpb_particle_mover_t)==16,     "particle_mover_t is 16 B" );
VPB_SA( sizeof(vpb_particle_injector_t)==48,  "particle_injector_t is 48 B" );
VPB_SA( sizeof(vpb_interpolator_t)==80,       "interpolator_t is 80 B (no USE_V8/V16)" );
VPB_SA( sizeof(vpb_hydro_t)==64,              "hydro_t is 64 B" );
VPB_SA( sizeof(vpb_accumulator_t)==48,        "accumulator_t is 48 B (no USE_V8/V16)" );
VPB_SA( sizeof(vpb_field_t)==80,              "field_t is 80 B" );
VPB_SA( offsetof(vpb_field_t,cbx)==16 && offsetof(vpb_field_t,tcax)==32 &&
        offsetof(vpb_field_t,jfx)==48 && offsetof(vpb_field_t,rhof)==60 &&
        offsetof(vpb_field_t,ematx)==64 && offsetof(vpb_field_t,cmat)==78, "field_t offsets" );
VPB_SA( sizeof(vpb_material_coefficient_t)==64, "material_coefficient_t is 64 B" );
VPB_SA( sizeof(vpb_species_t)==96,            "species_t is 96 B" );
VPB_SA( offsetof(vpb_species_t,p)==24 && offsetof(vpb_species_t,pm)==40 &&
        offsetof(vpb_species_t,partition)==64 && offsetof(vpb_species_t,g)==72, "species_t offsets" );
VPB_SA( sizeof(vpb_grid_t)==264,              "grid_t is 264 B" );
VPB_SA( offsetof(vpb_grid_t,nx)==56 && offsetof(vpb_grid_t,bc)==116 &&
        offsetof(vpb_grid_t,neighbor)==232 && offsetof(vpb_grid_t,rangel)==240, "grid_t offsets" );
VPB_SA( sizeof(vpb_field_advance_kernels_t)==17*sizeof(void*), "17 kernel-table entries" );
VPB_SA( sizeof(vpb_field_array_t)==24+17*sizeof(void*), "field_array_t" );
VPB_SA( sizeof(vpb_accumulator_array_t)==24 && sizeof(vpb_interpolator_array_t)==16, "sf arrays" );

#endif /* VPIC_B200_ABI_H */
