/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of boundary_p (F14) and accumulate_rhob (F15). */
#include "orc_common.h"

/* F15: src/species_advance/standard/rho_p.cc:125-171 (the scalar variant that
   is compiled in; its FMA/FNMS macros are separate multiply and add). */
void orc_accumulate_rhob( vpb_field_t * f, const vpb_particle_t * p, const vpb_grid_t * g, float qsp ) {
  float w0 = p->dx, w1 = p->dy, w2, w3, w4, w5, w6, w7, dz = p->dz;
  int v = p->i, x, y, z, sy = g->sy, sz = g->sz;
  w7 = (qsp*g->r8V)*p->w;
# define FMA( x,y,z) ((z)+(x)*(y))
# define FNMS(x,y,z) ((z)-(x)*(y))
  w6=FNMS(w0,w7,w7);
  w7=FMA( w0,w7,w7);
  w4=FNMS(w1,w6,w6); w5=FNMS(w1,w7,w7);
  w6=FMA( w1,w6,w6); w7=FMA( w1,w7,w7);
  w0=FNMS(dz,w4,w4); w1=FNMS(dz,w5,w5); w2=FNMS(dz,w6,w6); w3=FNMS(dz,w7,w7);
  w4=FMA( dz,w4,w4); w5=FMA( dz,w5,w5); w6=FMA( dz,w6,w6); w7=FMA( dz,w7,w7);
# undef FNMS
# undef FMA
  x = v; z = x/sz;
  if( z==1     ) w0 += w0, w1 += w1, w2 += w2, w3 += w3;
  if( z==g->nz ) w4 += w4, w5 += w5, w6 += w6, w7 += w7;
  x -= sz*z; y = x/sy;
  if( y==1     ) w0 += w0, w1 += w1, w4 += w4, w5 += w5;
  if( y==g->ny ) w2 += w2, w3 += w3, w6 += w6, w7 += w7;
  x -= sy*y;
  if( x==1     ) w0 += w0, w2 += w2, w4 += w4, w6 += w6;
  if( x==g->nx ) w1 += w1, w3 += w3, w5 += w5, w7 += w7;
  f[v      ].rhob += w0; f[v      +1].rhob += w1;
  f[v   +sy].rhob += w2; f[v   +sy+1].rhob += w3;
  f[v+sz   ].rhob += w4; f[v+sz   +1].rhob += w5;
  f[v+sz+sy].rhob += w6; f[v+sz+sy+1].rhob += w7;
}

#define ORC_MAX_SP 64

static void * orc_alloc( size_t bytes, size_t align ) {
  void * p = NULL;
  if( posix_memalign( &p, align, bytes ? bytes : align ) ) ORC_FAIL( "out of memory" );
  return p;
}

/* Exchange one pair of variable-length messages along axis X with the
   installed transport: counts first, then payloads (boundary_p.cc:292-316). */
static void orc_exchange_var( int X, int rem_lo, int rem_hi,
                              const void * slo, int64_t nlo, const void * shi, int64_t nhi,
                              void ** rlo, int64_t * got_lo, void ** rhi, int64_t * got_hi ) {
  int32_t cs[2] = { (int32_t)nlo, (int32_t)nhi }, cr[2] = { 0, 0 };
  int64_t sb[2] = { rem_lo ? 4 : 0, rem_hi ? 4 : 0 }, rb[2] = { sb[0], sb[1] };
  orc_exchange( orc_exchange_user, X, rem_lo ? &cs[0] : NULL, rem_hi ? &cs[1] : NULL,
                rem_lo ? &cr[0] : NULL, rem_hi ? &cr[1] : NULL, sb, rb );
  *got_lo = rem_lo ? cr[0] : 0; *got_hi = rem_hi ? cr[1] : 0;
  *rlo = orc_alloc( (size_t)( *got_lo ), 16 ); *rhi = orc_alloc( (size_t)( *got_hi ), 16 );
  sb[0] = rem_lo ? nlo : 0; sb[1] = rem_hi ? nhi : 0; rb[0] = *got_lo; rb[1] = *got_hi;
  orc_exchange( orc_exchange_user, X, rem_lo ? slo : NULL, rem_hi ? shi : NULL,
                rem_lo ? *rlo : NULL, rem_hi ? *rhi : NULL, sb, rb );
}

/* F14: src/boundary/boundary_p.cc:36-518 without custom handlers (pbc_list
   empty): absorb -> accumulate_rhob, remote neighbour -> injector on that
   face's message, holes back-filled from the end of the list (movers walked in
   reverse), received injectors appended in reverse order and finished with
   move_p on accumulator copy 0. */
void orc_boundary_p( vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa ) {
  orc_boundary_p_pbc( NULL, sp_list, fa, aa );
}

/* The same with the user handler list (boundary_p.cc:81-89, 236-253, 426-432):
   a particle that hits a face whose neighbor[] code is a handler id is given
   to that handler; the injectors the handlers return are injected first. */
void orc_boundary_p_pbc( vpb_particle_bc_t * pbc_list, vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa ) {
  if( !sp_list ) return;
  if( !fa || !aa || sp_list->g!=aa->g || fa->g!=aa->g ) ORC_FAIL( "Bad args" );
  static const int f2b[6] = { VPB_BOUNDARY(-1,0,0), VPB_BOUNDARY(0,-1,0), VPB_BOUNDARY(0,0,-1),
                              VPB_BOUNDARY( 1,0,0), VPB_BOUNDARY(0, 1,0), VPB_BOUNDARY(0,0, 1) };
  static const int axis[6] = { 0, 1, 2, 0, 1, 2 };
  static const float dir[6] = { 1, 1, 1, -1, -1, -1 };
  vpb_field_t * f = fa->f; vpb_grid_t * g = fa->g; vpb_accumulator_t * a0 = aa->a;
  const int64_t * neighbor = g->neighbor;
  const int64_t rangel = g->rangel, rangeh = g->rangeh, rangem = g->range[orc_world_size];
  int bc[6], shared[6]; int64_t range[6];
  for( int face=0; face<6; face++ ) {
    bc[face] = g->bc[f2b[face]];
    shared[face] = bc[face]>=0 && bc[face]<orc_world_size && bc[face]!=orc_world_rank;
    range[face] = shared[face] ? g->range[bc[face]] : 0;
  }
  int nm_tot = 0;
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) nm_tot += sp->nm;
  enum { ORC_MAX_PBC = 32 };
  vpb_particle_bc_func_t pbc_interact[ORC_MAX_PBC] = { 0 }; void * pbc_params[ORC_MAX_PBC] = { 0 };
  int nb = pbc_list ? (int)( -pbc_list->id-2 ) : 0;
  if( nb>ORC_MAX_PBC ) ORC_FAIL( "Update this to support more particle boundary conditions" );
  for( vpb_particle_bc_t * pbc=pbc_list; pbc; pbc=pbc->next ) {
    pbc_interact[-pbc->id-3] = pbc->interact; pbc_params[-pbc->id-3] = pbc->params;
  }
  vpb_particle_injector_t * ci = (vpb_particle_injector_t *)orc_alloc( sizeof(vpb_particle_injector_t)*(size_t)( nm_tot+1 ), 16 );
  int n_ci = 0;
  vpb_particle_injector_t * pi_send[6]; int n_send[6], n_recv[6];
  vpb_particle_injector_t * pi_recv[6];
  for( int face=0; face<6; face++ ) {
    pi_send[face] = shared[face] ? (vpb_particle_injector_t *)orc_alloc( sizeof(vpb_particle_injector_t)*(size_t)nm_tot, 16 ) : NULL;
    pi_recv[face] = NULL; n_send[face] = n_recv[face] = 0;
  }

  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    const float sp_q = sp->q; const int32_t sp_id = sp->id;
    vpb_particle_t * p0 = sp->p; int np = sp->np;
    vpb_particle_mover_t * pm = sp->pm + sp->nm - 1;
    for( int nm=sp->nm; nm; pm--, nm-- ) {
      int i = pm->i, voxel = p0[i].i, face = voxel & 7;
      voxel >>= 3; p0[i].i = voxel;
      int64_t nn = neighbor[ 6*(int64_t)voxel + face ];
      if( nn==VPB_ABSORB_PARTICLES ) orc_accumulate_rhob( f, p0+i, g, sp_q );
      else if( ( nn>=0 && nn<rangel ) || ( nn>rangeh && nn<=rangem ) ) {
        if( !shared[face] ) ORC_FAIL( "particle leaves through a face that has no remote neighbour" );
        vpb_particle_injector_t * pi = &pi_send[face][ n_send[face]++ ];
        pi->dx = p0[i].dx; pi->dy = p0[i].dy; pi->dz = p0[i].dz;
        pi->ux = p0[i].ux; pi->uy = p0[i].uy; pi->uz = p0[i].uz; pi->w = p0[i].w;
        pi->dispx = pm->dispx; pi->dispy = pm->dispy; pi->dispz = pm->dispz;
        (&pi->dx)[axis[face]] = dir[face];
        pi->i = (int32_t)( nn - range[face] );
        pi->sp_id = sp_id;
      } else if( -nn-3>=0 && -nn-3<nb ) {
        n_ci += pbc_interact[-nn-3]( pbc_params[-nn-3], sp, p0+i, pm, ci+n_ci, 1, face );
      } else fprintf( stderr, "oracle warning: unknown boundary interaction ... dropping particle (species=%s)\n", sp->name );
      np--;
      p0[i] = p0[np];
    }
    sp->np = np; sp->nm = 0;
  }

  /* exchange, one axis at a time (messages of different faces are independent) */
  for( int X=0; X<3; X++ ) {
    const int lo = X, hi = X+3;
    if( !shared[lo] && !shared[hi] ) continue;
    if( !orc_exchange ) ORC_FAIL( "remote neighbour but no transport installed" );
    void * rl, * rh; int64_t gl, gh;
    orc_exchange_var( X, shared[lo], shared[hi],
                      pi_send[lo], (int64_t)sizeof(vpb_particle_injector_t)*n_send[lo],
                      pi_send[hi], (int64_t)sizeof(vpb_particle_injector_t)*n_send[hi], &rl, &gl, &rh, &gh );
    pi_recv[lo] = (vpb_particle_injector_t *)rl; n_recv[lo] = (int)( gl/(int64_t)sizeof(vpb_particle_injector_t) );
    pi_recv[hi] = (vpb_particle_injector_t *)rh; n_recv[hi] = (int)( gh/(int64_t)sizeof(vpb_particle_injector_t) );
  }

  /* grow storage for the worst case (boundary_p.cc:318-406; shrinking is not restated) */
  int max_inj = n_ci;
  for( int face=0; face<6; face++ ) max_inj += n_recv[face];
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    int n = sp->np + max_inj;
    if( n>sp->max_np ) {
      n += 0.3125*n;
      vpb_particle_t * np_ = (vpb_particle_t *)orc_alloc( sizeof(vpb_particle_t)*(size_t)n, 128 );
      memcpy( np_, sp->p, sizeof(vpb_particle_t)*(size_t)sp->np );
      free( sp->p ); sp->p = np_; sp->max_np = n;
    }
    int nm = sp->nm + max_inj;
    if( nm>sp->max_nm ) {
      nm += 0.3125*nm;
      vpb_particle_mover_t * npm = (vpb_particle_mover_t *)orc_alloc( sizeof(vpb_particle_mover_t)*(size_t)nm, 128 );
      memcpy( npm, sp->pm, sizeof(vpb_particle_mover_t)*(size_t)sp->nm );
      free( sp->pm ); sp->pm = npm; sp->max_nm = nm;
    }
  }

  vpb_species_t * by_id[ORC_MAX_SP] = { 0 };
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    if( sp->id<0 || sp->id>=ORC_MAX_SP ) ORC_FAIL( "species id out of range" );
    by_id[sp->id] = sp;
  }
  for( int k=0; k<7; k++ ) {                             /* custom local injection first, then faces 0..5 */
    const int fc = k==0 ? 6 : k-1;
    if( fc<6 && !shared[fc] ) continue;
    const int cnt = fc==6 ? n_ci : n_recv[fc];
    const vpb_particle_injector_t * pi = ( fc==6 ? ci : pi_recv[fc] ) + cnt - 1;
    for( int n=cnt; n; pi--, n-- ) {
      vpb_species_t * sp = ( pi->sp_id>=0 && pi->sp_id<ORC_MAX_SP ) ? by_id[pi->sp_id] : NULL;
      if( !sp ) ORC_FAIL( "received an injector for an unknown species" );
      vpb_particle_t * p = sp->p; vpb_particle_mover_t * pm = sp->pm; int np = sp->np, nm = sp->nm;
      p[np].dx = pi->dx; p[np].dy = pi->dy; p[np].dz = pi->dz; p[np].i = pi->i;
      p[np].ux = pi->ux; p[np].uy = pi->uy; p[np].uz = pi->uz; p[np].w = pi->w;
      sp->np = np+1;
      pm[nm].dispx = pi->dispx; pm[nm].dispy = pi->dispy; pm[nm].dispz = pi->dispz; pm[nm].i = np;
      sp->nm = nm + orc_move_p( p, pm+nm, a0, g, sp->q );
    }
  }
  for( int face=0; face<6; face++ ) { free( pi_send[face] ); free( pi_recv[face] ); }
  free( ci );
}

/* A deterministic user handler for tests: the particle comes back from the
   wall with the momentum component along the face axis reversed and halved,
   the other two scaled by 3/4, and a quarter of its remaining displacement
   turned around; params counts the hits per face (int[6]). */
int orc_fixture_bounce_handler( void * params, vpb_species_t * sp, vpb_particle_t * p, vpb_particle_mover_t * pm,
                                vpb_particle_injector_t * pi, int max_pi, int face ) {
  (void)sp;
  if( max_pi<1 ) return 0;
  if( params ) ( (int *)params )[face]++;
  const int ax = face%3;
  float u[3] = { 0.75f*p->ux, 0.75f*p->uy, 0.75f*p->uz }, d[3] = { 0.25f*pm->dispx, 0.25f*pm->dispy, 0.25f*pm->dispz };
  u[ax] = -0.5f*(&p->ux)[ax];
  d[ax] = -d[ax];
  pi->dx = p->dx; pi->dy = p->dy; pi->dz = p->dz; pi->i = p->i;
  pi->ux = u[0]; pi->uy = u[1]; pi->uz = u[2]; pi->w = p->w;
  pi->dispx = d[0]; pi->dispy = d[1]; pi->dispz = d[2]; pi->sp_id = sp->id;
  return 1;
}
