/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of the field kernels: clear_jf (F8), synchronize_jf (F9),
 * advance_b (F10), advance_e (F11, vacuum and material), energy_f (F12), the
 * local boundary conditions of local.c and the face exchanges of remote.c. */
#include "orc_common.h"

typedef struct { int lo[3], hi[3]; } orc_box_t;

/* The reference's *_EDGE/_NODE/_FACE_LOOP(plane) index sets for the plane
   normal to axis X (remote.c:16-41): "YZ edge" = Y:1..nY, Z:1..nZ+1 (the
   Y-oriented edges), "ZY edge" = Y:1..nY+1, Z:1..nZ. */
enum { ORC_YZ_EDGE, ORC_ZY_EDGE, ORC_NODE, ORC_FACE };
static orc_box_t orc_plane( const vpb_grid_t * g, int X, int plane, int kind ) {
  const int n[3] = { g->nx, g->ny, g->nz };
  const int Y = (X+1)%3, Z = (X+2)%3;
  orc_box_t b;
  b.lo[X] = b.hi[X] = plane;
  b.lo[Y] = 1; b.lo[Z] = 1;
  b.hi[Y] = n[Y] + ( kind==ORC_ZY_EDGE || kind==ORC_NODE ? 1 : 0 );
  b.hi[Z] = n[Z] + ( kind==ORC_YZ_EDGE || kind==ORC_NODE ? 1 : 0 );
  return b;
}
#define ORC_BOX_LOOP(b) \
  for( int z=(b).lo[2]; z<=(b).hi[2]; z++ ) \
    for( int y=(b).lo[1]; y<=(b).hi[1]; y++ ) \
      for( int x=(b).lo[0]; x<=(b).hi[0]; x++ )
static inline int orc_box_count( orc_box_t b ) {
  return (b.hi[0]-b.lo[0]+1)*(b.hi[1]-b.lo[1]+1)*(b.hi[2]-b.lo[2]+1);
}

#define E_(f,c)   ((&(f)->ex  )[c])
#define CB_(f,c)  ((&(f)->cbx )[c])
#define TCA_(f,c) ((&(f)->tcax)[c])
#define JF_(f,c)  ((&(f)->jfx )[c])

static inline int orc_face_bc( const vpb_grid_t * g, int X, int side /*-1,+1*/ ) {
  int ijk[3] = {0,0,0}; ijk[X] = side;
  return g->bc[ VPB_BOUNDARY(ijk[0],ijk[1],ijk[2]) ];
}
static inline int orc_face_is_remote( const vpb_grid_t * g, int X, int side ) {
  int bc = orc_face_bc( g, X, side );
  return bc>=0 && bc<orc_world_size;
}

/* Move one pair of face messages along axis X.  Loop-back when no transport
   callback is installed (single rank talking to itself, grid_comm.c:20-23). */
static void orc_face_exchange( const vpb_grid_t * g, int X,
                               const float * slo, const float * shi, float * rlo, float * rhi,
                               int nfloat, int * got_lo, int * got_hi ) {
  int rem_lo = orc_face_is_remote( g, X, -1 ), rem_hi = orc_face_is_remote( g, X, +1 );
  *got_lo = rem_lo; *got_hi = rem_hi;
  if( !rem_lo && !rem_hi ) return;
  if( orc_exchange ) {
    int64_t sb[2] = { rem_lo ? 4*(int64_t)nfloat : 0, rem_hi ? 4*(int64_t)nfloat : 0 };
    int64_t rb[2] = { sb[0], sb[1] };
    orc_exchange( orc_exchange_user, X, rem_lo ? slo : NULL, rem_hi ? shi : NULL,
                  rem_lo ? rlo : NULL, rem_hi ? rhi : NULL, sb, rb );
    return;
  }
  if( orc_face_bc( g, X, -1 )!=orc_world_rank && rem_lo ) ORC_FAIL( "remote neighbour but no transport installed" );
  if( orc_face_bc( g, X, +1 )!=orc_world_rank && rem_hi ) ORC_FAIL( "remote neighbour but no transport installed" );
  /* what leaves through -X arrives on our own +X port and vice versa */
  if( rem_hi ) memcpy( rhi, slo, 4*(size_t)nfloat );
  if( rem_lo ) memcpy( rlo, shi, 4*(size_t)nfloat );
}

/* F8 */
void orc_clear_jf( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  for( int v=0; v<fa->g->nv; v++ ) fa->f[v].jfx = 0, fa->f[v].jfy = 0, fa->f[v].jfz = 0;
}

/* local.c:335-368 */
static void orc_local_adjust_jf( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    int bc = orc_face_bc( g, X, side );
    if( !orc_bc_is_local( bc ) ) continue;
    int face = side<0 ? 1 : n[X]+1, Y = (X+1)%3, Z = (X+2)%3;
    orc_box_t by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
    switch( bc ) {
    case VPB_PEC_FIELDS:
      ORC_BOX_LOOP(by) JF_( f+ORC_VOX(x,y,z), Y ) = 0;
      ORC_BOX_LOOP(bz) JF_( f+ORC_VOX(x,y,z), Z ) = 0;
      break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS: case VPB_ABSORB_FIELDS:
      ORC_BOX_LOOP(by) JF_( f+ORC_VOX(x,y,z), Y ) *= 2.;
      ORC_BOX_LOOP(bz) JF_( f+ORC_VOX(x,y,z), Z ) *= 2.;
      break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
  }
}

/* F9: remote.c:417-508.  Axis by axis; both messages of an axis are packed
   before either is applied. */
void orc_synchronize_jf( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  orc_local_adjust_jf( f, g );
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + n[Y]*(n[Z]+1) + n[Z]*(n[Y]+1);
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {          /* k=0: message out of -X (plane 1), k=1: out of +X (plane nX+1) */
      float * p = s[k]; int face = k==0 ? 1 : n[X]+1;
      orc_box_t by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
      *(p++) = d[X];
      ORC_BOX_LOOP(by) *(p++) = JF_( f+ORC_VOX(x,y,z), Y );
      ORC_BOX_LOOP(bz) *(p++) = JF_( f+ORC_VOX(x,y,z), Z );
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      /* r[0] came from the -X neighbour (it sent its plane nX+1, which is our
         plane 1); r[1] came from the +X neighbour (its plane 1 = our nX+1).
         The reference names receives by the REMOTE sending port
         (grid_comm.c:7-9), hence its "face = (i<0) ? nX+1 : 1" (remote.c:456). */
      if( !got[k] ) continue;
      const float * p = r[k]; int face = k==0 ? 1 : n[X]+1;
      float rw = *(p++), lw = rw + d[X];
      rw /= lw; lw = d[X]/lw; lw += lw; rw += rw;
      orc_box_t by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
      ORC_BOX_LOOP(by) { float * v = &JF_( f+ORC_VOX(x,y,z), Y ); *v = lw*(*v) + rw*(*(p++)); }
      ORC_BOX_LOOP(bz) { float * v = &JF_( f+ORC_VOX(x,y,z), Z ); *v = lw*(*v) + rw*(*(p++)); }
    }
    free( buf );
  }
}

/* local.c:267-296 */
static void orc_local_adjust_norm_b( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    int bc = orc_face_bc( g, X, side );
    if( !orc_bc_is_local( bc ) ) continue;
    int face = side<0 ? 1 : n[X]+1;
    switch( bc ) {
    case VPB_PEC_FIELDS: case VPB_PMC_FIELDS: case VPB_ABSORB_FIELDS: break;
    case VPB_SYMMETRIC_FIELDS: {
      orc_box_t b = orc_plane( g, X, face, ORC_FACE );
      ORC_BOX_LOOP(b) CB_( f+ORC_VOX(x,y,z), X ) = 0;
    } break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
  }
}

/* local.c:224-264 */
static void orc_local_adjust_tang_e( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    int bc = orc_face_bc( g, X, side );
    if( !orc_bc_is_local( bc ) ) continue;
    int face = side<0 ? 1 : n[X]+1, Y = (X+1)%3, Z = (X+2)%3;
    switch( bc ) {
    case VPB_PEC_FIELDS: {
      orc_box_t by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
      ORC_BOX_LOOP(by) { vpb_field_t * fs = f+ORC_VOX(x,y,z); E_(fs,Y) = 0; TCA_(fs,Y) = 0; }
      ORC_BOX_LOOP(bz) { vpb_field_t * fs = f+ORC_VOX(x,y,z); E_(fs,Z) = 0; TCA_(fs,Z) = 0; }
    } break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS: case VPB_ABSORB_FIELDS: break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
  }
}

/* local.c:50-122.  The absorbing branch is the 2nd-order implementation of a
   1st-order Higdon ABC; the ghost cb values it writes are state carried from
   step to step. */
static void orc_local_ghost_tang_b( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float cdt_d[3] = { g->cvac*g->dt*g->rdx, g->cvac*g->dt*g->rdy, g->cvac*g->dt*g->rdz };
  const float higend = ( nx>1 || ny>1 || nz>1 ) ? 1.03527618 : 1.;
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {
    int bc = orc_face_bc( g, X, side );
    if( !orc_bc_is_local( bc ) ) continue;
    const int Y = (X+1)%3, Z = (X+2)%3;
    int ghost = side<0 ? 0 : n[X]+1, face = side<0 ? 1 : n[X]+1;
    int off[3] = {0,0,0}; off[X] = side;      /* (i,j,k) */
    orc_box_t bY = orc_plane( g, X, ghost, ORC_ZY_EDGE );  /* cbY lives here */
    orc_box_t bZ = orc_plane( g, X, ghost, ORC_YZ_EDGE );  /* cbZ lives here */
#   define NB(x,y,z) ( f + ORC_VOX( (x)-off[0], (y)-off[1], (z)-off[2] ) )
    switch( bc ) {
    case VPB_PEC_FIELDS:
      ORC_BOX_LOOP(bY) CB_( f+ORC_VOX(x,y,z), Y ) =  CB_( NB(x,y,z), Y );
      ORC_BOX_LOOP(bZ) CB_( f+ORC_VOX(x,y,z), Z ) =  CB_( NB(x,y,z), Z );
      break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS:
      ORC_BOX_LOOP(bY) CB_( f+ORC_VOX(x,y,z), Y ) = -CB_( NB(x,y,z), Y );
      ORC_BOX_LOOP(bZ) CB_( f+ORC_VOX(x,y,z), Z ) = -CB_( NB(x,y,z), Z );
      break;
    case VPB_ABSORB_FIELDS: {
      float drive = cdt_d[X]*higend;
      float decay = (1-drive)/(1+drive);
      drive = 2*drive/(1+drive);
      ORC_BOX_LOOP(bY) {
        int c[3] = { x, y, z };
        vpb_field_t * fg = f + ORC_VOX(c[0],c[1],c[2]);
        vpb_field_t * fh = f + ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]);
        c[X] = face;
        float t1 = cdt_d[X]*( E_( f+ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]), Z ) - E_( f+ORC_VOX(c[0],c[1],c[2]), Z ) );
        t1 = side<0 ? t1 : -t1;
        c[X] = ghost;
        c[Z]++; float t2 = E_( f+ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]), X );
        c[Z]--; t2 = cdt_d[Z]*( t2 - E_(fh,X) );
        CB_(fg,Y) = decay*CB_(fg,Y) + drive*CB_(fh,Y) - t1 + t2;
      }
      ORC_BOX_LOOP(bZ) {
        int c[3] = { x, y, z };
        vpb_field_t * fg = f + ORC_VOX(c[0],c[1],c[2]);
        vpb_field_t * fh = f + ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]);
        c[X] = face;
        float t1 = cdt_d[X]*( E_( f+ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]), Y ) - E_( f+ORC_VOX(c[0],c[1],c[2]), Y ) );
        t1 = side<0 ? t1 : -t1;
        c[X] = ghost;
        c[Y]++; float t2 = E_( f+ORC_VOX(c[0]-off[0],c[1]-off[1],c[2]-off[2]), X );
        c[Y]--; t2 = cdt_d[Y]*( t2 - E_(fh,X) );
        CB_(fg,Z) = decay*CB_(fg,Z) + drive*CB_(fh,Z) + t1 - t2;
      }
    } break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
#   undef NB
  }
}

/* F10: advance_b_pipeline.{h,cc}.  cbx on (1:nx+1,1:ny,1:nz), cby on
   (1:nx,1:ny+1,1:nz), cbz on (1:nx,1:ny,1:nz+1). */
void orc_advance_b( vpb_field_array_t * fa, float frac ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const float px = (nx>1) ? frac*g->cvac*g->dt*g->rdx : 0;
  const float py = (ny>1) ? frac*g->cvac*g->dt*g->rdy : 0;
  const float pz = (nz>1) ? frac*g->cvac*g->dt*g->rdz : 0;
  for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
    vpb_field_t * f0 = f+ORC_VOX(x,y,z);
    const vpb_field_t * fx = f+ORC_VOX(x+1,y,z), * fy = f+ORC_VOX(x,y+1,z), * fz = f+ORC_VOX(x,y,z+1);
    if( y<=ny && z<=nz ) f0->cbx -= ( py*( fy->ez-f0->ez ) - pz*( fz->ey-f0->ey ) );
    if( z<=nz && x<=nx ) f0->cby -= ( pz*( fz->ex-f0->ex ) - px*( fx->ez-f0->ez ) );
    if( x<=nx && y<=ny ) f0->cbz -= ( px*( fx->ey-f0->ey ) - py*( fy->ex-f0->ex ) );
  }
  orc_local_adjust_norm_b( f, g );
}

/* remote.c:61-134: fill the tangential-B ghost planes from remote neighbours.
   A message out of port -X carries plane 1, out of +X plane nX; what arrives
   from the -X neighbour (its plane nX) lands in ghost plane 0, from the +X
   neighbour (its plane 1) in ghost plane nX+1. */
static void orc_remote_ghost_tang_b( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + n[Y]*(n[Z]+1) + n[Z]*(n[Y]+1);
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k]; int face = k==0 ? 1 : n[X];
      orc_box_t bY = orc_plane( g, X, face, ORC_ZY_EDGE ), bZ = orc_plane( g, X, face, ORC_YZ_EDGE );
      *(p++) = d[X];
      ORC_BOX_LOOP(bY) *(p++) = CB_( f+ORC_VOX(x,y,z), Y );
      ORC_BOX_LOOP(bZ) *(p++) = CB_( f+ORC_VOX(x,y,z), Z );
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      if( !got[k] ) continue;
      const float * p = r[k];
      int ghost = k==0 ? 0 : n[X]+1;
      int off[3] = {0,0,0}; off[X] = k==0 ? 1 : -1;   /* towards the interior */
      float lw = *(p++);
      float rw = (2.*d[X])/(lw+d[X]);
      lw = (lw-d[X])/(lw+d[X]);
      orc_box_t bY = orc_plane( g, X, ghost, ORC_ZY_EDGE ), bZ = orc_plane( g, X, ghost, ORC_YZ_EDGE );
      ORC_BOX_LOOP(bY) CB_( f+ORC_VOX(x,y,z), Y ) = rw*(*(p++)) + lw*CB_( f+ORC_VOX(x+off[0],y+off[1],z+off[2]), Y );
      ORC_BOX_LOOP(bZ) CB_( f+ORC_VOX(x,y,z), Z ) = rw*(*(p++)) + lw*CB_( f+ORC_VOX(x+off[0],y+off[1],z+off[2]), Z );
    }
    free( buf );
  }
}

/* F11: {vacuum_,}advance_e_pipeline.{h,cc}.  ex on (1:nx,1:ny+1,1:nz+1), ey on
   (1:nx+1,1:ny,1:nz+1), ez on (1:nx+1,1:ny+1,1:nz); the vacuum variant is what
   the reference installs when the material list has one entry (sfa.c:202-211). */
void orc_advance_e( vpb_field_array_t * fa, float frac ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  if( frac!=1 ) ORC_FAIL( "standard advance_e does not support frac != 1 yet" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * m = prm->mc;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const float damp = prm->damp;
  const float cj = g->dt/g->eps0;

  /* ghosts: remote data is packed from the pre-update cb, local ghosts use the
     pre-update e; both only write ghost cb, so doing them up front is
     equivalent to the reference's overlapped begin/end ordering. */
  orc_local_ghost_tang_b( f, g );
  orc_remote_ghost_tang_b( f, g );

  if( prm->n_mc==1 ) {
    const float decayx = m->decayx, drivex = m->drivex, decayy = m->decayy, drivey = m->drivey,
                decayz = m->decayz, drivez = m->drivez;
    const float px_muz = ((nx>1) ? (1+damp)*g->cvac*g->dt*g->rdx : 0)*m->rmuz;
    const float px_muy = ((nx>1) ? (1+damp)*g->cvac*g->dt*g->rdx : 0)*m->rmuy;
    const float py_mux = ((ny>1) ? (1+damp)*g->cvac*g->dt*g->rdy : 0)*m->rmux;
    const float py_muz = ((ny>1) ? (1+damp)*g->cvac*g->dt*g->rdy : 0)*m->rmuz;
    const float pz_muy = ((nz>1) ? (1+damp)*g->cvac*g->dt*g->rdz : 0)*m->rmuy;
    const float pz_mux = ((nz>1) ? (1+damp)*g->cvac*g->dt*g->rdz : 0)*m->rmux;
    for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z);
      const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
      if( x<=nx ) {
        f0->tcax = ( py_muz*( f0->cbz - fy->cbz ) - pz_muy*( f0->cby - fz->cby ) ) - damp*f0->tcax;
        f0->ex   = decayx*f0->ex + drivex*( f0->tcax - cj*f0->jfx );
      }
      if( y<=ny ) {
        f0->tcay = ( pz_mux*( f0->cbx - fz->cbx ) - px_muz*( f0->cbz - fx->cbz ) ) - damp*f0->tcay;
        f0->ey   = decayy*f0->ey + drivey*( f0->tcay - cj*f0->jfy );
      }
      if( z<=nz ) {
        f0->tcaz = ( px_muy*( f0->cby - fx->cby ) - py_mux*( f0->cbx - fy->cbx ) ) - damp*f0->tcaz;
        f0->ez   = decayz*f0->ez + drivez*( f0->tcaz - cj*f0->jfz );
      }
    }
  } else {
    const float px = (nx>1) ? (1+damp)*g->cvac*g->dt*g->rdx : 0;
    const float py = (ny>1) ? (1+damp)*g->cvac*g->dt*g->rdy : 0;
    const float pz = (nz>1) ? (1+damp)*g->cvac*g->dt*g->rdz : 0;
    for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z);
      const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
      if( x<=nx ) {
        f0->tcax = ( py*( f0->cbz*m[f0->fmatz].rmuz - fy->cbz*m[fy->fmatz].rmuz ) -
                     pz*( f0->cby*m[f0->fmaty].rmuy - fz->cby*m[fz->fmaty].rmuy ) ) - damp*f0->tcax;
        f0->ex   = m[f0->ematx].decayx*f0->ex + m[f0->ematx].drivex*( f0->tcax - cj*f0->jfx );
      }
      if( y<=ny ) {
        f0->tcay = ( pz*( f0->cbx*m[f0->fmatx].rmux - fz->cbx*m[fz->fmatx].rmux ) -
                     px*( f0->cbz*m[f0->fmatz].rmuz - fx->cbz*m[fx->fmatz].rmuz ) ) - damp*f0->tcay;
        f0->ey   = m[f0->ematy].decayy*f0->ey + m[f0->ematy].drivey*( f0->tcay - cj*f0->jfy );
      }
      if( z<=nz ) {
        f0->tcaz = ( px*( f0->cby*m[f0->fmaty].rmuy - fx->cby*m[fx->fmaty].rmuy ) -
                     py*( f0->cbx*m[f0->fmatx].rmux - fy->cbx*m[fy->fmatx].rmux ) ) - damp*f0->tcaz;
        f0->ez   = m[f0->ematz].decayz*f0->ez + m[f0->ematz].drivez*( f0->tcaz - cj*f0->jfz );
      }
    }
  }
  orc_local_adjust_tang_e( f, g );
}

/* F12: {vacuum_,}energy_f_pipeline.{h,c}.  Float products, double running sums
   per pipeline share (bulk in blocks of 16 voxels, then the host's remainder),
   times 0.5*eps0*dV.  Single-rank: no allreduce. */
extern int orc_energy_pipelines;
void orc_energy_f( double * en, const vpb_field_array_t * fa ) {
  if( !en || !fa ) ORC_FAIL( "Bad args" );
  const vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * m = prm->mc;
  const int nx = g->nx, ny = g->ny, nz = g->nz, P = orc_energy_pipelines;
  const int vacuum = prm->n_mc==1;
  const float qepsx = 0.25*m->epsx, qepsy = 0.25*m->epsy, qepsz = 0.25*m->epsz;
  const float hrmux = 0.50*m->rmux, hrmuy = 0.50*m->rmuy, hrmuz = 0.50*m->rmuz;
  double tot[6] = {0,0,0,0,0,0};
  const int ncell = nx*ny*nz;
  for( int r=0; r<=P; r++ ) {
    /* DISTRIBUTE_VOXELS(1,nx,1,ny,1,nz,16,r,P,...) (grid.h:318-333) */
    double t = (double)( ncell/16 )/(double)P;
    int i0 = 16*(int)( t*(double)r + 0.5 ), i1 = ncell;
    if( r<P ) i1 = 16*(int)( t*(double)(r+1) + 0.5 );
    double e[6] = {0,0,0,0,0,0};
    for( int i=i0; i<i1; i++ ) {
      int x = 1 + i%nx, y = 1 + (i/nx)%ny, z = 1 + i/(nx*ny);
      const vpb_field_t * f0 = f+ORC_VOX(x,y,z), * fx = f+ORC_VOX(x+1,y,z), * fy = f+ORC_VOX(x,y+1,z),
        * fz = f+ORC_VOX(x,y,z+1), * fyz = f+ORC_VOX(x,y+1,z+1), * fzx = f+ORC_VOX(x+1,y,z+1), * fxy = f+ORC_VOX(x+1,y+1,z);
      if( vacuum ) {
        e[0] += qepsx*( f0->ex*f0->ex + fy->ex*fy->ex + fz->ex*fz->ex + fyz->ex*fyz->ex );
        e[1] += qepsy*( f0->ey*f0->ey + fz->ey*fz->ey + fx->ey*fx->ey + fzx->ey*fzx->ey );
        e[2] += qepsz*( f0->ez*f0->ez + fx->ez*fx->ez + fy->ez*fy->ez + fxy->ez*fxy->ez );
        e[3] += hrmux*( f0->cbx*f0->cbx + fx->cbx*fx->cbx );
        e[4] += hrmuy*( f0->cby*f0->cby + fy->cby*fy->cby );
        e[5] += hrmuz*( f0->cbz*f0->cbz + fz->cbz*fz->cbz );
      } else {
        e[0] += 0.25*( m[ f0->ematx].epsx* f0->ex* f0->ex + m[ fy->ematx].epsx* fy->ex* fy->ex +
                       m[ fz->ematx].epsx* fz->ex* fz->ex + m[fyz->ematx].epsx*fyz->ex*fyz->ex );
        e[1] += 0.25*( m[ f0->ematy].epsy* f0->ey* f0->ey + m[ fz->ematy].epsy* fz->ey* fz->ey +
                       m[ fx->ematy].epsy* fx->ey* fx->ey + m[fzx->ematy].epsy*fzx->ey*fzx->ey );
        e[2] += 0.25*( m[ f0->ematz].epsz* f0->ez* f0->ez + m[ fx->ematz].epsz* fx->ez* fx->ez +
                       m[ fy->ematz].epsz* fy->ez* fy->ez + m[fxy->ematz].epsz*fxy->ez*fxy->ez );
        e[3] += 0.5 *( m[ f0->fmatx].rmux* f0->cbx* f0->cbx + m[ fx->fmatx].rmux* fx->cbx* fx->cbx );
        e[4] += 0.5 *( m[ f0->fmaty].rmuy* f0->cby* f0->cby + m[ fy->fmaty].rmuy* fy->cby* fy->cby );
        e[5] += 0.5 *( m[ f0->fmatz].rmuz* f0->cbz* f0->cbz + m[ fz->fmatz].rmuz* fz->cbz* fz->cbz );
      }
    }
    for( int c=0; c<6; c++ ) tot[c] += e[c];
  }
  double v0 = 0.5*g->eps0*g->dV;
  for( int c=0; c<6; c++ ) en[c] = tot[c]*v0;
}

/* sfa.c:55-151 (coefficient part only; the sanity warnings are host chatter). */
int orc_material_coefficients( vpb_material_coefficient_t * mc, int max_mc,
                               const vpb_material_t * m_list, const vpb_grid_t * g, float damp ) {
  (void)damp;
  int n_mc = 0;
  for( const vpb_material_t * m=m_list; m; m=m->next ) n_mc++;
  if( n_mc>max_mc ) ORC_FAIL( "material table too small" );
  for( const vpb_material_t * m=m_list; m; m=m->next ) {
    vpb_material_coefficient_t * c = mc + m->id;
    memset( c, 0, sizeof(*c) );
    float ax = ( m->sigmax*g->dt )/( m->epsx*g->eps0 );
    float ay = ( m->sigmay*g->dt )/( m->epsy*g->eps0 );
    float az = ( m->sigmaz*g->dt )/( m->epsz*g->eps0 );
    c->decayx = exp(-ax); c->decayy = exp(-ay); c->decayz = exp(-az);
    if( ax==0 ) c->drivex = 1./m->epsx; else if( c->decayx==0 ) c->drivex = 0;
    else c->drivex = 2.*exp(-0.5*ax)*sinh(0.5*ax)/(ax*m->epsx);
    if( ay==0 ) c->drivey = 1./m->epsy; else if( c->decayy==0 ) c->drivey = 0;
    else c->drivey = 2.*exp(-0.5*ay)*sinh(0.5*ay)/(ay*m->epsy);
    if( az==0 ) c->drivez = 1./m->epsz; else if( c->decayz==0 ) c->drivez = 0;
    else c->drivez = 2.*exp(-0.5*az)*sinh(0.5*az)/(az*m->epsz);
    c->rmux = 1./m->mux; c->rmuy = 1./m->muy; c->rmuz = 1./m->muz;
    c->nonconductive = ( ax==0 && ay==0 && az==0 ) ? 1. : 0.;
    c->epsx = m->epsx; c->epsy = m->epsy; c->epsz = m->epsz;
  }
  return n_mc;
}

/* ===========================================================================
 * N1: divergence cleaning family, charge-density synchronisation, shared-face
 * synchronisation, and the initialisation kernels compute_rhob /
 * compute_curl_b.  Reference: sfa_private.h:256-373 and the files cited at
 * each function.
 * ======================================================================== */

#define ORC_FOR_LOCAL_FACES( g, n, X, side, bc )                                  \
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ )                   \
    for( int bc = orc_face_bc( g, X, side ), once_ = orc_bc_is_local( bc ); once_; once_ = 0 )

/* sfa.c:239-245 */
void orc_clear_rhof( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  for( int v=0; v<fa->g->nv; v++ ) fa->f[v].rhof = 0;
}

/* rho_p.cc:22-112: trilinear charge of every particle onto the 8 nodes of its
   cell, in particle order (separate multiply and add, as compiled). */
void orc_accumulate_rho_p( vpb_field_array_t * fa, const vpb_species_t * sp ) {
  if( !fa || !sp || fa->g!=sp->g ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_particle_t * p = sp->p;
  const float q_8V = sp->q*sp->g->r8V;
  const int np = sp->np, sy = sp->g->sy, sz = sp->g->sz;
  for( int n=0; n<np; n++ ) {
    float w0 = p[n].dx, w1 = p[n].dy, w2, w3, w4, w5, w6, w7, dz = p[n].dz;
    const int v = p[n].i;
    w7 = p[n].w*q_8V;
    w6 = w7 - w0*w7; w7 = w7 + w0*w7;
    w4 = w6 - w1*w6; w5 = w7 - w1*w7;
    w6 = w6 + w1*w6; w7 = w7 + w1*w7;
    w0 = w4 - dz*w4; w1 = w5 - dz*w5; w2 = w6 - dz*w6; w3 = w7 - dz*w7;
    w4 = w4 + dz*w4; w5 = w5 + dz*w5; w6 = w6 + dz*w6; w7 = w7 + dz*w7;
    f[v      ].rhof += w0; f[v      +1].rhof += w1;
    f[v   +sy].rhof += w2; f[v   +sy+1].rhof += w3;
    f[v+sz   ].rhof += w4; f[v+sz   +1].rhof += w5;
    f[v+sz+sy].rhof += w6; f[v+sz+sy+1].rhof += w7;
  }
}

/* local.c:376-445 */
static void orc_local_adjust_rho( vpb_field_t * f, const vpb_grid_t * g, int which /*0 rhof, 1 rhob*/ ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  ORC_FOR_LOCAL_FACES( g, n, X, side, bc ) {
    orc_box_t b = orc_plane( g, X, side<0 ? 1 : n[X]+1, ORC_NODE );
    switch( bc ) {
    case VPB_PEC_FIELDS:
      ORC_BOX_LOOP(b) { if( which ) f[ORC_VOX(x,y,z)].rhob = 0; else f[ORC_VOX(x,y,z)].rhof = 0; }
      break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS: case VPB_ABSORB_FIELDS:
      if( !which ) ORC_BOX_LOOP(b) f[ORC_VOX(x,y,z)].rhof *= 2;
      break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
  }
}

/* remote.c:534-624: axis by axis, node planes 1 and nX+1 of rhof and rhob are
   folded with the neighbour's (rhof summed, rhob averaged, for equal cells). */
void orc_synchronize_rho( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  orc_local_adjust_rho( f, g, 0 );
  orc_local_adjust_rho( f, g, 1 );
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + 2*(n[Y]+1)*(n[Z]+1);
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k];
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X]+1, ORC_NODE );
      *(p++) = d[X];
      ORC_BOX_LOOP(b) { const vpb_field_t * f0 = f+ORC_VOX(x,y,z); *(p++) = f0->rhof; *(p++) = f0->rhob; }
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      if( !got[k] ) continue;
      const float * p = r[k];
      float hrw = *(p++), hlw = hrw + d[X], lw, rw;
      hrw /= hlw; hlw = d[X]/hlw; lw = hlw + hlw; rw = hrw + hrw;
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X]+1, ORC_NODE );
      ORC_BOX_LOOP(b) {
        vpb_field_t * f0 = f+ORC_VOX(x,y,z);
        f0->rhof =  lw*f0->rhof +  rw*(*(p++));
        f0->rhob = hlw*f0->rhob + hrw*(*(p++));
      }
    }
    free( buf );
  }
}

/* local.c:128-179: normal-E (and its tca) ghost nodes of locally handled faces */
static void orc_local_ghost_norm_e( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  ORC_FOR_LOCAL_FACES( g, n, X, side, bc ) {
    int off[3] = {0,0,0}; off[X] = side;
    orc_box_t b = orc_plane( g, X, side<0 ? 0 : n[X]+1, ORC_NODE );
    ORC_BOX_LOOP(b) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z);
      const vpb_field_t * f1 = f+ORC_VOX(x-off[0],y-off[1],z-off[2]), * f2 = f+ORC_VOX(x-2*off[0],y-2*off[1],z-2*off[2]);
      switch( bc ) {
      case VPB_PEC_FIELDS:       E_(f0,X) =  E_(f1,X); TCA_(f0,X) =  TCA_(f1,X); break;
      case VPB_SYMMETRIC_FIELDS:
      case VPB_PMC_FIELDS:       E_(f0,X) = -E_(f1,X); TCA_(f0,X) = -TCA_(f1,X); break;
      case VPB_ABSORB_FIELDS:    E_(f0,X) = 2*E_(f1,X) - E_(f2,X); TCA_(f0,X) = 2*TCA_(f1,X) - TCA_(f2,X); break;
      default: ORC_FAIL( "Bad boundary condition encountered." );
      }
    }
  }
}

/* remote.c:137-206: the -X port carries eX of node plane 1, the +X port of
   plane nX; arrivals land in ghost planes 0 / nX+1 (tca is not exchanged). */
static void orc_remote_ghost_norm_e( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + (n[Y]+1)*(n[Z]+1);
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k];
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X], ORC_NODE );
      *(p++) = d[X];
      ORC_BOX_LOOP(b) *(p++) = E_( f+ORC_VOX(x,y,z), X );
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      if( !got[k] ) continue;
      const float * p = r[k];
      int off[3] = {0,0,0}; off[X] = k==0 ? 1 : -1;   /* towards the interior */
      float lw = *(p++);
      float rw = (2.*d[X])/(lw+d[X]);
      lw = (lw-d[X])/(lw+d[X]);
      orc_box_t b = orc_plane( g, X, k==0 ? 0 : n[X]+1, ORC_NODE );
      ORC_BOX_LOOP(b) E_( f+ORC_VOX(x,y,z), X ) = rw*(*(p++)) + lw*E_( f+ORC_VOX(x+off[0],y+off[1],z+off[2]), X );
    }
    free( buf );
  }
}

/* compute_div_e_err / compute_rhob and their vacuum variants
   ({vacuum_,}compute_{div_e_err,rhob}_pipeline.{h,c}): every node
   (1:nx+1,1:ny+1,1:nz+1) once; the variants differ in how the constants are
   folded, which changes the rounding and is kept. */
static void orc_div_e_stencil( vpb_field_array_t * fa, int want_rhob ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * m = prm->mc;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  orc_local_ghost_norm_e( f, g );
  orc_remote_ghost_norm_e( f, g );
  const int vacuum = prm->n_mc==1;
  const float nc = m->nonconductive;
  float px, py, pz; const float cj = 1./g->eps0;
  if( want_rhob ) {
    if( vacuum ) {
      px = (nx>1) ? g->eps0*m->epsx*g->rdx : 0; py = (ny>1) ? g->eps0*m->epsy*g->rdy : 0; pz = (nz>1) ? g->eps0*m->epsz*g->rdz : 0;
    } else {
      px = (nx>1) ? g->eps0*g->rdx : 0; py = (ny>1) ? g->eps0*g->rdy : 0; pz = (nz>1) ? g->eps0*g->rdz : 0;
    }
  } else {
    if( vacuum ) {
      px = ((nx>1) ? g->rdx : 0)*m->epsx; py = ((ny>1) ? g->rdy : 0)*m->epsy; pz = ((nz>1) ? g->rdz : 0)*m->epsz;
    } else {
      px = (nx>1) ? g->rdx : 0; py = (ny>1) ? g->rdy : 0; pz = (nz>1) ? g->rdz : 0;
    }
  }
  for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
    vpb_field_t * f0 = f+ORC_VOX(x,y,z);
    const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
    if( vacuum ) {
      if( want_rhob ) f0->rhob = nc*( px*( f0->ex - fx->ex ) + py*( f0->ey - fy->ey ) + pz*( f0->ez - fz->ez ) - f0->rhof );
      else f0->div_e_err = nc*( px*( f0->ex - fx->ex ) + py*( f0->ey - fy->ey ) + pz*( f0->ez - fz->ez ) - cj*( f0->rhof + f0->rhob ) );
    } else {
      if( want_rhob ) f0->rhob = m[f0->nmat].nonconductive *
        ( px*( m[f0->ematx].epsx*f0->ex - m[fx->ematx].epsx*fx->ex ) +
          py*( m[f0->ematy].epsy*f0->ey - m[fy->ematy].epsy*fy->ey ) +
          pz*( m[f0->ematz].epsz*f0->ez - m[fz->ematz].epsz*fz->ez ) - f0->rhof );
      else f0->div_e_err = m[f0->nmat].nonconductive *
        ( px*( m[f0->ematx].epsx*f0->ex - m[fx->ematx].epsx*fx->ex ) +
          py*( m[f0->ematy].epsy*f0->ey - m[fy->ematy].epsy*fy->ey ) +
          pz*( m[f0->ematz].epsz*f0->ez - m[fz->ematz].epsz*fz->ez ) - cj*( f0->rhof + f0->rhob ) );
    }
  }
}

/* compute_div_e_err_pipeline.c:43-190 + local.c:298-333 */
void orc_compute_div_e_err( vpb_field_array_t * fa ) {
  orc_div_e_stencil( fa, 0 );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  ORC_FOR_LOCAL_FACES( g, n, X, side, bc ) {
    orc_box_t b = orc_plane( g, X, side<0 ? 1 : n[X]+1, ORC_NODE );
    switch( bc ) {
    case VPB_PEC_FIELDS: case VPB_ABSORB_FIELDS: ORC_BOX_LOOP(b) f[ORC_VOX(x,y,z)].div_e_err = 0; break;
    case VPB_SYMMETRIC_FIELDS: case VPB_PMC_FIELDS: break;
    default: ORC_FAIL( "Bad boundary condition encountered." );
    }
  }
}

/* compute_rhob_pipeline.c:43-175 + local_adjust_rhob */
void orc_compute_rhob( vpb_field_array_t * fa ) {
  orc_div_e_stencil( fa, 1 );
  orc_local_adjust_rho( fa->f, fa->g, 1 );
}

/* Sum of doubles across ranks through the face transport's sibling: tests
   install orc_allsum for multi-rank runs; single rank is the identity. */
void (*orc_allsum)( double * v, int n ) = 0;
void orc_set_allsum( void (*fn)( double *, int ) ) { orc_allsum = fn; }

/* compute_rms_div_e_err_pipeline.c:53-166: interior nodes as FLOAT squares
   added to a double, boundary nodes with weights 1/2, 1/4, 1/8 as double
   products.  Order: the exterior first (host), then the pipelines' shares. */
double orc_compute_rms_div_e_err( const vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  const vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  double err = 0;
# define ORC_SQ(x,y,z,w) do { const vpb_field_t * f0 = f+ORC_VOX(x,y,z); err += (w)*(double)f0->div_e_err*(double)f0->div_e_err; } while(0)
  for( int y=2; y<=ny; y++ ) for( int z=2; z<=nz; z++ ) { ORC_SQ(1,y,z,0.5); ORC_SQ(nx+1,y,z,0.5); }
  for( int z=2; z<=nz; z++ ) for( int x=2; x<=nx; x++ ) { ORC_SQ(x,1,z,0.5); ORC_SQ(x,ny+1,z,0.5); }
  for( int x=2; x<=nx; x++ ) for( int y=2; y<=ny; y++ ) { ORC_SQ(x,y,1,0.5); ORC_SQ(x,y,nz+1,0.5); }
  for( int x=2; x<=nx; x++ ) { ORC_SQ(x,1,1,0.25); ORC_SQ(x,ny+1,1,0.25); ORC_SQ(x,1,nz+1,0.25); ORC_SQ(x,ny+1,nz+1,0.25); }
  for( int y=2; y<=ny; y++ ) { ORC_SQ(1,y,1,0.25); ORC_SQ(1,y,nz+1,0.25); ORC_SQ(nx+1,y,1,0.25); ORC_SQ(nx+1,y,nz+1,0.25); }
  for( int z=2; z<=nz; z++ ) { ORC_SQ(1,1,z,0.25); ORC_SQ(nx+1,1,z,0.25); ORC_SQ(1,ny+1,z,0.25); ORC_SQ(nx+1,ny+1,z,0.25); }
  ORC_SQ(1,1,1,0.125); ORC_SQ(nx+1,1,1,0.125); ORC_SQ(1,ny+1,1,0.125); ORC_SQ(nx+1,ny+1,1,0.125);
  ORC_SQ(1,1,nz+1,0.125); ORC_SQ(nx+1,1,nz+1,0.125); ORC_SQ(1,ny+1,nz+1,0.125); ORC_SQ(nx+1,ny+1,nz+1,0.125);
# undef ORC_SQ
  /* interior: the pipelines' shares (blocks of 16 voxels) and the host's
     straggler share, each summed on its own, added in pipeline order */
  {
    const int bx = nx-1, by = ny-1, bz = nz-1, P = orc_energy_pipelines;
    const int nvox = ( bx>0 && by>0 && bz>0 ) ? bx*by*bz : 0;
    for( int r=0; r<=P; r++ ) {
      double t = (double)( nvox/16 )/(double)P;
      int i0 = 16*(int)( t*(double)r + 0.5 ), i1 = nvox;
      if( r<P ) i1 = 16*(int)( t*(double)(r+1) + 0.5 );
      double part = 0;
      for( int i=i0; i<i1; i++ ) {
        const vpb_field_t * f0 = f+ORC_VOX( 2 + i%bx, 2 + (i/bx)%by, 2 + i/(bx*by) );
        part += f0->div_e_err*f0->div_e_err;
      }
      err += part;
    }
  }
  double local[2] = { err*g->dV, ( g->nx*g->ny*g->nz )*g->dV };
  if( orc_allsum ) orc_allsum( local, 2 );
  return g->eps0*sqrt( local[0]/local[1] );
}

/* {vacuum_,}clean_div_e_pipeline.{h,c} + local_adjust_tang_e: Marder pass on
   ex (1:nx,1:ny+1,1:nz+1), ey (1:nx+1,1:ny,1:nz+1), ez (1:nx+1,1:ny+1,1:nz). */
void orc_clean_div_e( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * m = prm->mc;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const int vacuum = prm->n_mc==1;
  const float _rdx = (nx>1) ? g->rdx : 0, _rdy = (ny>1) ? g->rdy : 0, _rdz = (nz>1) ? g->rdz : 0;
  const float alphadt = 0.3888889/( _rdx*_rdx + _rdy*_rdy + _rdz*_rdz );
  float px, py, pz;
  if( vacuum ) { px = (alphadt*_rdx)*m->drivex; py = (alphadt*_rdy)*m->drivey; pz = (alphadt*_rdz)*m->drivez; }
  else         { px = alphadt*_rdx; py = alphadt*_rdy; pz = alphadt*_rdz; }
  for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
    vpb_field_t * f0 = f+ORC_VOX(x,y,z);
    const vpb_field_t * fx = f+ORC_VOX(x+1,y,z), * fy = f+ORC_VOX(x,y+1,z), * fz = f+ORC_VOX(x,y,z+1);
    if( vacuum ) {
      if( x<=nx ) f0->ex += px*( fx->div_e_err - f0->div_e_err );
      if( y<=ny ) f0->ey += py*( fy->div_e_err - f0->div_e_err );
      if( z<=nz ) f0->ez += pz*( fz->div_e_err - f0->div_e_err );
    } else {
      if( x<=nx ) f0->ex += m[f0->ematx].drivex*px*( fx->div_e_err - f0->div_e_err );
      if( y<=ny ) f0->ey += m[f0->ematy].drivey*py*( fy->div_e_err - f0->div_e_err );
      if( z<=nz ) f0->ez += m[f0->ematz].drivez*pz*( fz->div_e_err - f0->div_e_err );
    }
  }
  orc_local_adjust_tang_e( f, g );
}

/* compute_div_b_err_pipeline.cc:7-53: cells (1:nx,1:ny,1:nz) */
void orc_compute_div_b_err( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const float px = (nx>1) ? g->rdx : 0, py = (ny>1) ? g->rdy : 0, pz = (nz>1) ? g->rdz : 0;
  for( int z=1; z<=nz; z++ ) for( int y=1; y<=ny; y++ ) for( int x=1; x<=nx; x++ ) {
    vpb_field_t * f0 = f+ORC_VOX(x,y,z);
    const vpb_field_t * fx = f+ORC_VOX(x+1,y,z), * fy = f+ORC_VOX(x,y+1,z), * fz = f+ORC_VOX(x,y,z+1);
    f0->div_b_err = px*( fx->cbx - f0->cbx ) + py*( fy->cby - f0->cby ) + pz*( fz->cbz - f0->cbz );
  }
}

/* compute_rms_div_b_err_pipeline.c:7-103 */
double orc_compute_rms_div_b_err( const vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  const vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz, P = orc_energy_pipelines;
  const int nvox = nx*ny*nz;
  double err = 0;
  for( int r=0; r<=P; r++ ) {
    double t = (double)( nvox/16 )/(double)P;
    int i0 = 16*(int)( t*(double)r + 0.5 ), i1 = nvox;
    if( r<P ) i1 = 16*(int)( t*(double)(r+1) + 0.5 );
    double part = 0;
    for( int i=i0; i<i1; i++ ) {
      const vpb_field_t * f0 = f+ORC_VOX( 1 + i%nx, 1 + (i/nx)%ny, 1 + i/(nx*ny) );
      part += f0->div_b_err*f0->div_b_err;
    }
    err += part;
  }
  double local[2] = { err*g->dV, ( g->nx*g->ny*g->nz )*g->dV };
  if( orc_allsum ) orc_allsum( local, 2 );
  return g->eps0*sqrt( local[0]/local[1] );
}

/* local.c:181-222 */
static void orc_local_ghost_div_b( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  ORC_FOR_LOCAL_FACES( g, n, X, side, bc ) {
    int off[3] = {0,0,0}; off[X] = side;
    orc_box_t b = orc_plane( g, X, side<0 ? 0 : n[X]+1, ORC_FACE );
    ORC_BOX_LOOP(b) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z); const vpb_field_t * f1 = f+ORC_VOX(x-off[0],y-off[1],z-off[2]);
      switch( bc ) {
      case VPB_PEC_FIELDS:       f0->div_b_err =  f1->div_b_err; break;
      case VPB_SYMMETRIC_FIELDS:
      case VPB_PMC_FIELDS:       f0->div_b_err = -f1->div_b_err; break;
      case VPB_ABSORB_FIELDS:    f0->div_b_err = 0; break;
      default: ORC_FAIL( "Bad boundary condition encountered." );
      }
    }
  }
}

/* remote.c:208-296 */
static void orc_remote_ghost_div_b( vpb_field_t * f, const vpb_grid_t * g ) {
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + n[Y]*n[Z];
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k];
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X], ORC_FACE );
      *(p++) = d[X];
      ORC_BOX_LOOP(b) *(p++) = f[ORC_VOX(x,y,z)].div_b_err;
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      if( !got[k] ) continue;
      const float * p = r[k];
      int off[3] = {0,0,0}; off[X] = k==0 ? 1 : -1;
      float lw = *(p++);
      float rw = (2.*d[X])/(lw+d[X]);
      lw = (lw-d[X])/(lw+d[X]);
      orc_box_t b = orc_plane( g, X, k==0 ? 0 : n[X]+1, ORC_FACE );
      ORC_BOX_LOOP(b) f[ORC_VOX(x,y,z)].div_b_err = rw*(*(p++)) + lw*f[ORC_VOX(x+off[0],y+off[1],z+off[2])].div_b_err;
    }
    free( buf );
  }
}

/* clean_div_b_pipeline.cc:17-307 + local_adjust_norm_b: cbx on
   (1:nx+1,1:ny,1:nz), cby on (1:nx,1:ny+1,1:nz), cbz on (1:nx,1:ny,1:nz+1). */
void orc_clean_div_b( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  float px = (nx>1) ? g->rdx : 0, py = (ny>1) ? g->rdy : 0, pz = (nz>1) ? g->rdz : 0;
  const float alphadt = 0.3888889/( px*px + py*py + pz*pz );
  px *= alphadt; py *= alphadt; pz *= alphadt;
  orc_local_ghost_div_b( f, g );
  orc_remote_ghost_div_b( f, g );
  for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
    vpb_field_t * f0 = f+ORC_VOX(x,y,z);
    const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
    if( y<=ny && z<=nz ) f0->cbx += px*( f0->div_b_err - fx->div_b_err );
    if( z<=nz && x<=nx ) f0->cby += py*( f0->div_b_err - fy->div_b_err );
    if( x<=nx && y<=ny ) f0->cbz += pz*( f0->div_b_err - fz->div_b_err );
  }
  orc_local_adjust_norm_b( f, g );
}

/* {vacuum_,}compute_curl_b_pipeline.{h,cc}: tca = curl( cB/mu_r ) on the same
   edge sets as advance_e, with the same ghost / adjust steps around it. */
void orc_compute_curl_b( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const vpb_sfa_params_t * prm = (const vpb_sfa_params_t *)fa->params;
  const vpb_material_coefficient_t * m = prm->mc;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  orc_local_ghost_tang_b( f, g );
  orc_remote_ghost_tang_b( f, g );
  if( prm->n_mc==1 ) {
    const float px_muz = ((nx>1) ? g->cvac*g->dt*g->rdx : 0)*m->rmuz;
    const float px_muy = ((nx>1) ? g->cvac*g->dt*g->rdx : 0)*m->rmuy;
    const float py_mux = ((ny>1) ? g->cvac*g->dt*g->rdy : 0)*m->rmux;
    const float py_muz = ((ny>1) ? g->cvac*g->dt*g->rdy : 0)*m->rmuz;
    const float pz_muy = ((nz>1) ? g->cvac*g->dt*g->rdz : 0)*m->rmuy;
    const float pz_mux = ((nz>1) ? g->cvac*g->dt*g->rdz : 0)*m->rmux;
    for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z);
      const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
      if( x<=nx ) f0->tcax = ( py_muz*( f0->cbz - fy->cbz ) - pz_muy*( f0->cby - fz->cby ) );
      if( y<=ny ) f0->tcay = ( pz_mux*( f0->cbx - fz->cbx ) - px_muz*( f0->cbz - fx->cbz ) );
      if( z<=nz ) f0->tcaz = ( px_muy*( f0->cby - fx->cby ) - py_mux*( f0->cbx - fy->cbx ) );
    }
  } else {
    const float px = (nx>1) ? g->cvac*g->dt*g->rdx : 0;
    const float py = (ny>1) ? g->cvac*g->dt*g->rdy : 0;
    const float pz = (nz>1) ? g->cvac*g->dt*g->rdz : 0;
    for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
      vpb_field_t * f0 = f+ORC_VOX(x,y,z);
      const vpb_field_t * fx = f+ORC_VOX(x-1,y,z), * fy = f+ORC_VOX(x,y-1,z), * fz = f+ORC_VOX(x,y,z-1);
      if( x<=nx ) f0->tcax = ( py*( f0->cbz*m[f0->fmatz].rmuz - fy->cbz*m[fy->fmatz].rmuz ) -
                               pz*( f0->cby*m[f0->fmaty].rmuy - fz->cby*m[fz->fmaty].rmuy ) );
      if( y<=ny ) f0->tcay = ( pz*( f0->cbx*m[f0->fmatx].rmux - fz->cbx*m[fz->fmatx].rmux ) -
                               px*( f0->cbz*m[f0->fmatz].rmuz - fx->cbz*m[fx->fmatz].rmuz ) );
      if( z<=nz ) f0->tcaz = ( px*( f0->cby*m[f0->fmaty].rmuy - fx->cby*m[fx->fmaty].rmuy ) -
                               py*( f0->cbx*m[f0->fmatx].rmux - fy->cbx*m[fy->fmatx].rmux ) );
    }
  }
  orc_local_adjust_tang_e( f, g );
}

/* remote.c:298-415: average the doubly owned face values (cbX on faces, eY/eZ
   and tcaY/tcaZ on edges of planes 1 and nX+1) with the neighbour's, axis by
   axis, and return the summed squared mismatch (cb and e only). */
double orc_synchronize_tang_e_norm_b( vpb_field_array_t * fa ) {
  if( !fa ) ORC_FAIL( "Bad args" );
  vpb_field_t * f = fa->f; const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  double err = 0;
  orc_local_adjust_tang_e( f, g );
  orc_local_adjust_norm_b( f, g );
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 2*n[Y]*(n[Z]+1) + 2*n[Z]*(n[Y]+1) + n[Y]*n[Z];
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k]; int face = k==0 ? 1 : n[X]+1;
      orc_box_t bf = orc_plane( g, X, face, ORC_FACE ), by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
      ORC_BOX_LOOP(bf) *(p++) = CB_( f+ORC_VOX(x,y,z), X );
      ORC_BOX_LOOP(by) { const vpb_field_t * f0 = f+ORC_VOX(x,y,z); *(p++) = E_(f0,Y); *(p++) = TCA_(f0,Y); }
      ORC_BOX_LOOP(bz) { const vpb_field_t * f0 = f+ORC_VOX(x,y,z); *(p++) = E_(f0,Z); *(p++) = TCA_(f0,Z); }
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    /* END_RECV(-1,0,0) comes first in the reference and names the message by
       the SENDER's port (see orc_synchronize_jf): it is the +X neighbour's
       plane 1 and updates our face nX+1 ("face = (i+j+k)<0 ? nX+1 : 1"). */
    for( int k=1; k>=0; k-- ) {          /* face nX+1 first: fixes the order of the double sum */
      if( !got[k] ) continue;
      const float * p = r[k]; int face = k==0 ? 1 : n[X]+1;
      double w1, w2;
      orc_box_t bf = orc_plane( g, X, face, ORC_FACE ), by = orc_plane( g, X, face, ORC_YZ_EDGE ), bz = orc_plane( g, X, face, ORC_ZY_EDGE );
      ORC_BOX_LOOP(bf) { vpb_field_t * f0 = f+ORC_VOX(x,y,z); w1 = *(p++); w2 = CB_(f0,X); CB_(f0,X) = 0.5*( w1+w2 ); err += (w1-w2)*(w1-w2); }
      ORC_BOX_LOOP(by) { vpb_field_t * f0 = f+ORC_VOX(x,y,z);
        w1 = *(p++); w2 = E_(f0,Y);   E_(f0,Y)   = 0.5*( w1+w2 ); err += (w1-w2)*(w1-w2);
        w1 = *(p++); w2 = TCA_(f0,Y); TCA_(f0,Y) = 0.5*( w1+w2 ); }
      ORC_BOX_LOOP(bz) { vpb_field_t * f0 = f+ORC_VOX(x,y,z);
        w1 = *(p++); w2 = E_(f0,Z);   E_(f0,Z)   = 0.5*( w1+w2 ); err += (w1-w2)*(w1-w2);
        w1 = *(p++); w2 = TCA_(f0,Z); TCA_(f0,Z) = 0.5*( w1+w2 ); }
    }
    free( buf );
  }
  if( orc_allsum ) orc_allsum( &err, 1 );
  return err;
}

/* ---- hydro moments (8(f) N3) ------------------------------------------------ */

/* hydro_array.c:52-56 */
void orc_clear_hydro_array( vpb_hydro_array_t * ha ) {
  if( !ha ) ORC_FAIL( "NULL hydro array" );
  memset( ha->h, 0, sizeof(vpb_hydro_t)*(size_t)ha->g->nv );
}

/* hydro_p.c:26-167.  One serial loop over the particles: half a Boris step to
   time-centre the momentum, then the 14 moments spread trilinearly over the
   eight nodes of the cell.  sqrt() of the float argument is the DOUBLE sqrt
   (:91), and the (1./3.), 0.4 constants make :96 a double expression. */
void orc_accumulate_hydro_p( vpb_hydro_array_t * ha, const vpb_species_t * sp, const vpb_interpolator_array_t * ia ) {
  if( !ha || !sp || !ia || ha->g!=sp->g || ha->g!=ia->g ) ORC_FAIL( "Bad args" );
  vpb_hydro_t * h = ha->h; const vpb_particle_t * p = sp->p; const vpb_interpolator_t * f = ia->i;
  const vpb_grid_t * g = sp->g;
  const float c = g->cvac, qsp = sp->q, mspc = sp->m*c;
  const float qdt_2mc = (qsp*g->dt)/(2*mspc), qdt_4mc2 = qdt_2mc/(2*c), r8V = g->r8V;
  const int sx = 1, sy = g->nx+2, sz = (g->nx+2)*(g->ny+2);
  const int off[8] = { 0, sx, sy, sy+sx, sz, sz+sx, sz+sy, sz+sy+sx };
  for( int n=0; n<sp->np; n++ ) {
    float dx = p[n].dx, dy = p[n].dy, dz = p[n].dz; const int i = p[n].i;
    float ux = p[n].ux, uy = p[n].uy, uz = p[n].uz; const float w = p[n].w;
    float w0, w1, w2, w3, w4, w5, w6, w7, vx, vy, vz, ke_mc;
    ux += qdt_2mc*( ( f[i].ex + dy*f[i].dexdy ) + dz*( f[i].dexdz + dy*f[i].d2exdydz ) );
    uy += qdt_2mc*( ( f[i].ey + dz*f[i].deydz ) + dx*( f[i].deydx + dz*f[i].d2eydzdx ) );
    uz += qdt_2mc*( ( f[i].ez + dx*f[i].dezdx ) + dy*( f[i].dezdy + dx*f[i].d2ezdxdy ) );
    w5 = f[i].cbx + dx*f[i].dcbxdx;
    w6 = f[i].cby + dy*f[i].dcbydy;
    w7 = f[i].cbz + dz*f[i].dcbzdz;
    ke_mc = ux*ux + uy*uy + uz*uz;
    vz = (float)sqrt( (double)( 1+ke_mc ) );
    ke_mc *= c/( vz+1 );
    vz = c/vz;
    w0 = qdt_4mc2*vz;
    w1 = w5*w5 + w6*w6 + w7*w7;
    w2 = w0*w0*w1;
    w3 = (float)( (double)w0*( 1 + ( 1./3. )*(double)w2*( 1 + 0.4*(double)w2 ) ) );
    w4 = w3/( 1 + w1*w3*w3 ); w4 += w4;
    w0 = ux + w3*( uy*w7 - uz*w6 );
    w1 = uy + w3*( uz*w5 - ux*w7 );
    w2 = uz + w3*( ux*w6 - uy*w5 );
    ux += w4*( w1*w7 - w2*w6 );
    uy += w4*( w2*w5 - w0*w7 );
    uz += w4*( w0*w6 - w1*w5 );
    vx = ux*vz; vy = uy*vz; vz *= uz;
    w0 = r8V*w; dx *= w0; w1 = w0+dx; w0 -= dx;
    w3 = 1+dy; w2 = w0*w3; w3 *= w1; dy = 1-dy; w0 *= dy; w1 *= dy;
    w7 = 1+dz; w4 = w0*w7; w5 = w1*w7; w6 = w2*w7; w7 *= w3;
    dz = 1-dz; w0 *= dz; w1 *= dz; w2 *= dz; w3 *= dz;
    const float wn[8] = { w0, w1, w2, w3, w4, w5, w6, w7 };
    for( int k=0; k<8; k++ ) {
      vpb_hydro_t * hk = h + i + off[k];
      float t = qsp*wn[k];
      hk->jx += t*vx; hk->jy += t*vy; hk->jz += t*vz; hk->rho += t;
      t = mspc*wn[k];
      const float px = t*ux, py = t*uy, pz = t*uz;
      hk->px += px; hk->py += py; hk->pz += pz; hk->ke += t*ke_mc;
      hk->txx += px*vx; hk->tyy += py*vy; hk->tzz += pz*vz;
      hk->tyz += py*vz; hk->tzx += pz*vx; hk->txy += px*vy;
    }
  }
}

/* hydro_array.c:71-222: double the moments on locally handled faces (half
   cells), then the twice-weighted sum with each remote (or periodic-to-self)
   neighbour's copy of the shared node plane, axis by axis. */
void orc_synchronize_hydro_array( vpb_hydro_array_t * ha ) {
  if( !ha ) ORC_FAIL( "NULL hydro array" );
  vpb_hydro_t * h = ha->h; const vpb_grid_t * g = ha->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz; const int n[3] = { nx, ny, nz };
  const float d[3] = { g->dx, g->dy, g->dz };
  for( int side=-1; side<=1; side+=2 ) for( int X=0; X<3; X++ ) {          /* ADJUST_HYDRO order: -x,-y,-z,+x,+y,+z */
    if( !orc_bc_is_local( orc_face_bc( g, X, side ) ) ) continue;
    orc_box_t b = orc_plane( g, X, side<0 ? 1 : n[X]+1, ORC_NODE );
    ORC_BOX_LOOP(b) { float * v = &h[ORC_VOX(x,y,z)].jx; for( int c=0; c<14; c++ ) v[c] *= 2; }
  }
  for( int X=0; X<3; X++ ) {
    const int Y = (X+1)%3, Z = (X+2)%3;
    int nf = 1 + 14*(n[Y]+1)*(n[Z]+1);
    float * buf = (float *)malloc( 4*sizeof(float)*(size_t)nf );
    float * s[2] = { buf, buf+nf }, * r[2] = { buf+2*nf, buf+3*nf };
    for( int k=0; k<2; k++ ) {
      float * p = s[k];
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X]+1, ORC_NODE );
      *(p++) = d[X];
      ORC_BOX_LOOP(b) { const float * v = &h[ORC_VOX(x,y,z)].jx; for( int c=0; c<14; c++ ) *(p++) = v[c]; }
    }
    int got[2];
    orc_face_exchange( g, X, s[0], s[1], r[0], r[1], nf, &got[0], &got[1] );
    for( int k=0; k<2; k++ ) {
      if( !got[k] ) continue;
      const float * p = r[k];
      float rw = *(p++), lw = rw + d[X];
      rw /= lw; lw = d[X]/lw; lw += lw; rw += rw;
      orc_box_t b = orc_plane( g, X, k==0 ? 1 : n[X]+1, ORC_NODE );
      ORC_BOX_LOOP(b) { float * v = &h[ORC_VOX(x,y,z)].jx; for( int c=0; c<14; c++ ) v[c] = lw*v[c] + rw*(*(p++)); }
    }
    free( buf );
  }
}
