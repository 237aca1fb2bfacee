/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of the particle kernels: advance_p (F1), move_p (F2),
 * sort_p (F3), energy_p (F13), center_p/uncenter_p (N2). */
#include "orc_common.h"

/* Deposit one streak (midpoint m, half-displacement d, charge q, correction
   v5) into the 12 slots of accumulator a.  Operation order follows
   advance_p_pipeline.cc:183-205 / move_p.cc:276-296 exactly. */
static inline void orc_deposit( float * a, float q,
                                float dx, float dy, float dz,
                                float ux, float uy, float uz, float v5 ) {
  float v0, v1, v2, v3, v4;
# define ACC(uX,dY,dZ,o)             \
  v4 = q*uX;  v1 = v4*dY;            \
  v0 = v4-v1; v1 += v4;              \
  v4 = 1.0f+dZ; v2 = v0*v4; v3 = v1*v4; \
  v4 = 1.0f-dZ; v0 *= v4;   v1 *= v4;   \
  v0 += v5; v1 -= v5; v2 -= v5; v3 += v5; \
  a[o+0] += v0; a[o+1] += v1; a[o+2] += v2; a[o+3] += v3
  ACC( ux, dy, dz, 0 );
  ACC( uy, dz, dx, 4 );
  ACC( uz, dx, dy, 8 );
# undef ACC
}

/* F2: move_p.cc:216-353 (the scalar variant). */
int orc_move_p( vpb_particle_t * p0, vpb_particle_mover_t * pm, vpb_accumulator_t * a0,
                const vpb_grid_t * g, float qsp ) {
  vpb_particle_t * p = p0 + pm->i;
  float q = qsp*p->w;
  for(;;) {
    float s_midx = p->dx, s_midy = p->dy, s_midz = p->dz;
    float s_dispx = pm->dispx, s_dispy = pm->dispy, s_dispz = pm->dispz;
    float s_dir[3];
    s_dir[0] = (s_dispx>0.0f) ? 1.0f : -1.0f;
    s_dir[1] = (s_dispy>0.0f) ? 1.0f : -1.0f;
    s_dir[2] = (s_dispz>0.0f) ? 1.0f : -1.0f;
    /* twice the fractional distance to each candidate face */
    float v0 = (s_dispx==0.0f) ? 3.4e38f : (s_dir[0]-s_midx)/s_dispx;
    float v1 = (s_dispy==0.0f) ? 3.4e38f : (s_dir[1]-s_midy)/s_dispy;
    float v2 = (s_dispz==0.0f) ? 3.4e38f : (s_dir[2]-s_midz)/s_dispz;
    float v3 = 2.0f; int axis = 3;
    if( v0<v3 ) { v3 = v0; axis = 0; }
    if( v1<v3 ) { v3 = v1; axis = 1; }
    if( v2<v3 ) { v3 = v2; axis = 2; }
    v3 *= 0.5f;
    s_dispx *= v3; s_dispy *= v3; s_dispz *= v3;
    s_midx += s_dispx; s_midy += s_dispy; s_midz += s_dispz;
    /* NOTE the double constant (1./3.): product in float, promoted, rounded
       back once (move_p.cc:274). */
    float v5 = q*s_dispx*s_dispy*s_dispz*(1./3.);
    orc_deposit( (float *)( a0 + p->i ), q, s_midx, s_midy, s_midz, s_dispx, s_dispy, s_dispz, v5 );
    pm->dispx -= s_dispx; pm->dispy -= s_dispy; pm->dispz -= s_dispz;
    p->dx += s_dispx+s_dispx; p->dy += s_dispy+s_dispy; p->dz += s_dispz+s_dispz;
    if( axis==3 ) break;
    float dir = s_dir[axis];
    (&(p->dx))[axis] = dir;               /* exactly on the face */
    int face = axis; if( dir>0 ) face += 3;
    int64_t neighbor = g->neighbor[ 6*(int64_t)p->i + face ];
    if( neighbor==VPB_REFLECT_PARTICLES ) {
      (&(p->ux))[axis]     = -(&(p->ux))[axis];
      (&(pm->dispx))[axis] = -(&(pm->dispx))[axis];
      continue;
    }
    if( neighbor<g->rangel || neighbor>g->rangeh ) {
      p->i = 8*p->i + face;               /* hand over to boundary_p */
      return 1;
    }
    p->i = (int32_t)( neighbor - g->rangel );
    (&(p->dx))[axis] = -dir;
  }
  return 0;
}

/* One pipeline's share of advance_p (advance_p_pipeline.cc:20-237). */
static void orc_advance_p_range( vpb_species_t * sp, vpb_accumulator_t * a_all,
                                 vpb_accumulator_t * a0, const vpb_interpolator_t * f0,
                                 int first, int n, vpb_particle_mover_t * pm, int max_nm,
                                 int * nm_out, int * nignored_out ) {
  const vpb_grid_t * g = sp->g;
  const float qdt_2mc = (sp->q*g->dt)/(2*sp->m*g->cvac);
  const float cdt_dx  = g->cvac*g->dt*g->rdx;
  const float cdt_dy  = g->cvac*g->dt*g->rdy;
  const float cdt_dz  = g->cvac*g->dt*g->rdz;
  const float qsp = sp->q;
  const float one = 1.0f, one_third = 1.0/3.0, two_fifteenths = 2.0/15.0;
  int nm = 0, nignored = 0;
  (void)a_all;
  vpb_particle_t * p = sp->p + first;
  for( ; n; n--, p++ ) {
    float dx = p->dx, dy = p->dy, dz = p->dz; int ii = p->i;
    const vpb_interpolator_t * f = f0 + ii;
    float hax = qdt_2mc*( ( f->ex + dy*f->dexdy ) + dz*( f->dexdz + dy*f->d2exdydz ) );
    float hay = qdt_2mc*( ( f->ey + dz*f->deydz ) + dx*( f->deydx + dz*f->d2eydzdx ) );
    float haz = qdt_2mc*( ( f->ez + dx*f->dezdx ) + dy*( f->dezdy + dx*f->d2ezdxdy ) );
    float cbx = f->cbx + dx*f->dcbxdx;
    float cby = f->cby + dy*f->dcbydy;
    float cbz = f->cbz + dz*f->dcbzdz;
    float ux = p->ux, uy = p->uy, uz = p->uz, q = p->w;
    float v0, v1, v2, v3, v4, v5;
    ux += hax; uy += hay; uz += haz;
    v0 = qdt_2mc/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
    v1 = cbx*cbx + ( cby*cby + cbz*cbz );
    v2 = ( v0*v0 )*v1;
    v3 = v0*( one + v2*( one_third + v2*two_fifteenths ) );
    v4 = v3/( one + v1*( v3*v3 ) );
    v4 += v4;
    v0 = ux + v3*( uy*cbz - uz*cby );
    v1 = uy + v3*( uz*cbx - ux*cbz );
    v2 = uz + v3*( ux*cby - uy*cbx );
    ux += v4*( v1*cbz - v2*cby );
    uy += v4*( v2*cbx - v0*cbz );
    uz += v4*( v0*cby - v1*cbx );
    ux += hax; uy += hay; uz += haz;
    p->ux = ux; p->uy = uy; p->uz = uz;
    v0 = one/sqrtf( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
    ux *= cdt_dx; uy *= cdt_dy; uz *= cdt_dz;
    ux *= v0; uy *= v0; uz *= v0;
    v0 = dx + ux; v1 = dy + uy; v2 = dz + uz;      /* streak midpoint */
    v3 = v0 + ux; v4 = v1 + uy; v5 = v2 + uz;      /* new position    */
    if(  v3<=one &&  v4<=one &&  v5<=one && -v3<=one && -v4<=one && -v5<=one ) {
      q *= qsp;
      p->dx = v3; p->dy = v4; p->dz = v5;
      v5 = q*ux*uy*uz*one_third;
      orc_deposit( (float *)( a0 + ii ), q, v0, v1, v2, ux, uy, uz, v5 );
    } else {
      vpb_particle_mover_t local_pm;
      local_pm.dispx = ux; local_pm.dispy = uy; local_pm.dispz = uz;
      local_pm.i = (int32_t)( p - sp->p );
      if( orc_move_p( sp->p, &local_pm, a0, g, qsp ) ) {
        if( nm<max_nm ) pm[nm++] = local_pm; else nignored++;
      }
    }
  }
  *nm_out = nm; *nignored_out = nignored;
}

/* F1: advance_p_pipeline.cc:245-314.  Pipelines r=0..P-1 take DISTRIBUTE(np,16)
   shares into accumulator copy 1+r, the host takes the np%16 stragglers into
   copy 0; mover segments follow DISTRIBUTE(max_nm-(np&15),8) and are then
   compacted in pipeline order. */
void orc_advance_p( vpb_species_t * sp, vpb_accumulator_array_t * aa, const vpb_interpolator_array_t * ia ) {
  if( !sp || !aa || !ia || sp->g!=aa->g || sp->g!=ia->g ) ORC_FAIL( "Bad args" );
  const int P = aa->n_pipeline;
  sp->nm = 0;
  int seg_off[VPB_MAX_PIPELINE+1], seg_nm[VPB_MAX_PIPELINE+1];
  if( P>VPB_MAX_PIPELINE ) ORC_FAIL( "too many pipelines" );
  int max_nm_tot = sp->max_nm - ( sp->np&15 ); if( max_nm_tot<0 ) max_nm_tot = 0;
  for( int r=0; r<=P; r++ ) {
    int first, n, moff, mmax, nign;
    if( P==0 ) { first = 0; n = sp->np; moff = 0; mmax = sp->max_nm; }
    else {
      orc_distribute( sp->np, 16, r, P, &first, &n );
      orc_distribute( max_nm_tot, 8, r, P, &moff, &mmax );
      if( r==P ) mmax = sp->max_nm - moff;
    }
    vpb_accumulator_t * a0 = aa->a + ( r==P ? 0 : (size_t)(1+r)*aa->stride );
    orc_advance_p_range( sp, aa->a, a0, ia->i, first, n, sp->pm + moff, mmax, &seg_nm[r], &nign );
    seg_off[r] = moff;
    if( nign ) fprintf( stderr, "oracle warning: pipeline %d ran out of storage for %d movers\n", r, nign );
    if( P==0 ) break;
  }
  for( int r=0; r<=P; r++ ) {
    if( sp->nm!=seg_off[r] ) memmove( sp->pm + sp->nm, sp->pm + seg_off[r], sizeof(vpb_particle_mover_t)*(size_t)seg_nm[r] );
    sp->nm += seg_nm[r];
    if( P==0 ) break;
  }
}

/* F3: sort_p_pipeline.c:143-212 (subsort) + :335-363.  Stable counting sort by
   voxel; partition[v] = index of the first particle of voxel v for
   v in [vl,vh+1], 0 below vl, np above vh (entry nv is left untouched like the
   reference). */
void orc_sort_p( vpb_species_t * sp ) {
  if( !sp ) ORC_FAIL( "Bad args" );
  const vpb_grid_t * g = sp->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz, nv = g->nv, np = sp->np;
  const int vl = ORC_VOX(1,1,1), vh = ORC_VOX(nx,ny,nz);
  sp->last_sorted = g->step;
  int * next = (int *)calloc( (size_t)nv+1, sizeof(int) );
  vpb_particle_t * aux = (vpb_particle_t *)malloc( sizeof(vpb_particle_t)*(size_t)(np>0?np:1) );
  if( !next || !aux ) ORC_FAIL( "out of memory" );
  for( int i=0; i<np; i++ ) next[ sp->p[i].i ]++;
  int sum = 0;
  for( int v=vl; v<=vh; v++ ) { int c = next[v]; next[v] = sum; sp->partition[v] = sum; sum += c; }
  sp->partition[vh+1] = sum;
  for( int i=0; i<np; i++ ) { int j = next[ sp->p[i].i ]++; aux[j] = sp->p[i]; }
  for( int v=0; v<vl; v++ ) sp->partition[v] = 0;
  for( int v=vh+1; v<nv; v++ ) sp->partition[v] = np;
  memcpy( sp->p, aux, sizeof(vpb_particle_t)*(size_t)np );
  free( aux ); free( next );
}

/* F13: energy_p_pipeline.cc:17-115.  Per-pipeline double partial sums added in
   pipeline order, then times c^2. */
double orc_energy_p( const vpb_species_t * sp, const vpb_interpolator_array_t * ia ) {
  if( !sp || !ia || sp->g!=ia->g ) ORC_FAIL( "Bad args" );
  const vpb_grid_t * g = sp->g;
  const float qdt_2mc = (sp->q*g->dt)/(2*sp->m*g->cvac);
  const float msp = sp->m, one = 1.0f;
  const vpb_interpolator_t * f = ia->i;
  const vpb_particle_t * p = sp->p;
  /* n_pipeline is a process-wide setting in the reference; the accumulator
     array carries it, the species does not, so the caller passes it through
     orc_set_energy_pipelines (default 1). */
  extern int orc_energy_pipelines;
  const int P = orc_energy_pipelines;
  double local = 0.0;
  for( int r=0; r<=P; r++ ) {
    int n0, n1; double en = 0.0;
    orc_distribute( sp->np, 16, r, P, &n0, &n1 ); n1 += n0;
    for( int n=n0; n<n1; n++ ) {
      float dx = p[n].dx, dy = p[n].dy, dz = p[n].dz; int i = p[n].i;
      float v0 = p[n].ux + qdt_2mc*( ( f[i].ex + dy*f[i].dexdy ) + dz*( f[i].dexdz + dy*f[i].d2exdydz ) );
      float v1 = p[n].uy + qdt_2mc*( ( f[i].ey + dz*f[i].deydz ) + dx*( f[i].deydx + dz*f[i].d2eydzdx ) );
      float v2 = p[n].uz + qdt_2mc*( ( f[i].ez + dx*f[i].dezdx ) + dy*( f[i].dezdy + dx*f[i].d2ezdxdy ) );
      v0 = v0*v0 + v1*v1 + v2*v2;
      v0 = (msp*p[n].w)*( v0/( one + sqrtf( one + v0 ) ) );
      en += (double)v0;
    }
    local += en;
  }
  return local*( (double)g->cvac*(double)g->cvac );
}
int orc_energy_pipelines = 1;
void orc_set_energy_pipelines( int P ) { orc_energy_pipelines = P<1 ? 1 : P; }

/* N2: center_p / uncenter_p (center_p_pipeline.cc / uncenter_p_pipeline.cc
   :16-98): half E kick + half Boris rotation, forward or backward. */
static void orc_half_step( vpb_species_t * sp, const vpb_interpolator_array_t * ia, int uncenter ) {
  if( !sp || !ia || sp->g!=ia->g ) ORC_FAIL( "Bad args" );
  const vpb_grid_t * g = sp->g;
  const float args_qdt_2mc = (sp->q*g->dt)/(2*sp->m*g->cvac);
  const float qdt_2mc = uncenter ? -args_qdt_2mc : args_qdt_2mc;
  const float qdt_4mc = uncenter ? -0.5*args_qdt_2mc : 0.5*args_qdt_2mc;
  const float one = 1.0f, one_third = 1.0/3.0, two_fifteenths = 2.0/15.0;
  vpb_particle_t * p = sp->p;
  for( int n=sp->np; n; n--, p++ ) {
    float dx = p->dx, dy = p->dy, dz = p->dz;
    const vpb_interpolator_t * f = ia->i + p->i;
    float hax = qdt_2mc*( ( f->ex + dy*f->dexdy ) + dz*( f->dexdz + dy*f->d2exdydz ) );
    float hay = qdt_2mc*( ( f->ey + dz*f->deydz ) + dx*( f->deydx + dz*f->d2eydzdx ) );
    float haz = qdt_2mc*( ( f->ez + dx*f->dezdx ) + dy*( f->dezdy + dx*f->d2ezdxdy ) );
    float cbx = f->cbx + dx*f->dcbxdx, cby = f->cby + dy*f->dcbydy, cbz = f->cbz + dz*f->dcbzdz;
    float ux = p->ux, uy = p->uy, uz = p->uz, v0, v1, v2, v3, v4;
    if( !uncenter ) { ux += hax; uy += hay; uz += haz; }
    v0 = qdt_4mc/(float)sqrt( one + ( ux*ux + ( uy*uy + uz*uz ) ) );
    v1 = cbx*cbx + ( cby*cby + cbz*cbz );
    v2 = ( v0*v0 )*v1;
    v3 = v0*( one + v2*( one_third + v2*two_fifteenths ) );
    v4 = v3/( one + v1*( v3*v3 ) );
    v4 += v4;
    v0 = ux + v3*( uy*cbz - uz*cby );
    v1 = uy + v3*( uz*cbx - ux*cbz );
    v2 = uz + v3*( ux*cby - uy*cbx );
    ux += v4*( v1*cbz - v2*cby );
    uy += v4*( v2*cbx - v0*cbz );
    uz += v4*( v0*cby - v1*cbx );
    if( uncenter ) { ux += hax; uy += hay; uz += haz; }
    p->ux = ux; p->uy = uy; p->uz = uz;
  }
}
void orc_uncenter_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia ) { orc_half_step( sp, ia, 1 ); }
void orc_center_p( vpb_species_t * sp, const vpb_interpolator_array_t * ia ) { orc_half_step( sp, ia, 0 ); }

/* src/collision/langevin.c:12-24 + pipeline/langevin_pipeline.c:20-42, 44-112 (single pipeline: the whole
   array in order).  The reference draws frandn(rng) per component; here the deviates are an input. */
void orc_apply_langevin( vpb_langevin_t * l, const float * normals ) {
  vpb_species_t * sp = l->sp;
  if( l->interval<1 || ( sp->g->step % l->interval ) ) return;
  float nudt = l->nu * (float)l->interval * sp->g->dt;
  float decay = exp( -nudt );
  float drive = sqrt( ( -expm1( -2*nudt )*l->kT )/( sp->m*sp->g->cvac ) );
  vpb_particle_t * p = sp->p;
  for( int i=0; i<sp->np; i++ ) {
    p[i].ux = decay*p[i].ux + drive*( normals ? normals[3*i+0] : 0.f );
    p[i].uy = decay*p[i].uy + drive*( normals ? normals[3*i+1] : 0.f );
    p[i].uz = decay*p[i].uz + drive*( normals ? normals[3*i+2] : 0.f );
  }
}
