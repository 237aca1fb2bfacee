/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of the species<->field interface kernels (F4-F7). */
#include "orc_common.h"

int orc_world_rank = 0, orc_world_size = 1;
orc_exchange_fn orc_exchange = NULL; void * orc_exchange_user = NULL;

void orc_set_world( int rank, int size ) { orc_world_rank = rank; orc_world_size = size; }
void orc_set_exchange( orc_exchange_fn fn, void * user ) { orc_exchange = fn; orc_exchange_user = user; }

/* F7: interpolator_array_pipeline.cc:66-112.  The parenthesisation of the
   four-point combinations is part of the contract (bit-exact). */
void orc_load_interpolator_array( vpb_interpolator_array_t * ia, const vpb_field_array_t * fa ) {
  if( !ia || !fa || ia->g!=fa->g ) ORC_FAIL( "Bad args" );
  const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const vpb_field_t * f = fa->f;
  const float fourth = 0.25f, half = 0.5f;
  for( int z=1; z<=nz; z++ ) for( int y=1; y<=ny; y++ ) for( int x=1; x<=nx; x++ ) {
    vpb_interpolator_t * pi = ia->i + ORC_VOX(x,y,z);
    const vpb_field_t * f0  = f + ORC_VOX(x,  y,  z  );
    const vpb_field_t * fx  = f + ORC_VOX(x+1,y,  z  );
    const vpb_field_t * fy  = f + ORC_VOX(x,  y+1,z  );
    const vpb_field_t * fz  = f + ORC_VOX(x,  y,  z+1);
    const vpb_field_t * fyz = f + ORC_VOX(x,  y+1,z+1);
    const vpb_field_t * fzx = f + ORC_VOX(x+1,y,  z+1);
    const vpb_field_t * fxy = f + ORC_VOX(x+1,y+1,z  );
    float w0, w1, w2, w3;
    w0 = f0->ex; w1 = fy->ex; w2 = fz->ex; w3 = fyz->ex;
    pi->ex       = fourth*( (w3 + w0) + (w1 + w2) );
    pi->dexdy    = fourth*( (w3 - w0) + (w1 - w2) );
    pi->dexdz    = fourth*( (w3 - w0) - (w1 - w2) );
    pi->d2exdydz = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = f0->ey; w1 = fz->ey; w2 = fx->ey; w3 = fzx->ey;
    pi->ey       = fourth*( (w3 + w0) + (w1 + w2) );
    pi->deydz    = fourth*( (w3 - w0) + (w1 - w2) );
    pi->deydx    = fourth*( (w3 - w0) - (w1 - w2) );
    pi->d2eydzdx = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = f0->ez; w1 = fx->ez; w2 = fy->ez; w3 = fxy->ez;
    pi->ez       = fourth*( (w3 + w0) + (w1 + w2) );
    pi->dezdx    = fourth*( (w3 - w0) + (w1 - w2) );
    pi->dezdy    = fourth*( (w3 - w0) - (w1 - w2) );
    pi->d2ezdxdy = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = f0->cbx; w1 = fx->cbx; pi->cbx = half*( w1 + w0 ); pi->dcbxdx = half*( w1 - w0 );
    w0 = f0->cby; w1 = fy->cby; pi->cby = half*( w1 + w0 ); pi->dcbydy = half*( w1 - w0 );
    w0 = f0->cbz; w1 = fz->cbz; pi->cbz = half*( w1 + w0 ); pi->dcbzdz = half*( w1 - w0 );
  }
}

/* F4: clear_accumulators_pipeline.c:40-61 -- every copy, voxels
   [VOXEL(1,1,1) rounded down to even, VOXEL(nx,ny,nz)] rounded up to even count. */
void orc_clear_accumulator_array( vpb_accumulator_array_t * aa ) {
  if( !aa ) ORC_FAIL( "Bad args" );
  const int nx = aa->g->nx, ny = aa->g->ny, nz = aa->g->nz;
  int i0 = ( ORC_VOX(1,1,1)/2 )*2;
  int n  = ( ( ( ORC_VOX(nx,ny,nz) - i0 + 1 ) + 1 )/2 )*2;
  for( int c=0; c<=aa->n_pipeline; c++ )
    memset( aa->a + (size_t)c*aa->stride + i0, 0, sizeof(vpb_accumulator_t)*(size_t)n );
}

/* F5: reduce_accumulators_pipeline.cc:86-102 -- fixed association tree
   ((a0+a1)+(a2+a3))+... over the private copies, same voxel range as F4. */
void orc_reduce_accumulator_array( vpb_accumulator_array_t * aa ) {
  if( !aa ) ORC_FAIL( "Bad args" );
  const int nx = aa->g->nx, ny = aa->g->ny, nz = aa->g->nz;
  int i0 = ( ORC_VOX(1,1,1)/2 )*2;
  int n  = ( ( ( ORC_VOX(nx,ny,nz) - i0 + 1 ) + 1 )/2 )*2;
  int nr = aa->n_pipeline;
  if( nr>9 ) ORC_FAIL( "oracle supports at most 9 pipeline copies" );
  size_t sr = 12*(size_t)aa->stride;
  float * a = (float *)( aa->a + i0 );
  for( size_t k=0; k<12*(size_t)n; k++ ) {
    const float * b = a + k + sr;  /* b[r*sr] = copy r+1 */
    float v = a[k];
#   define B(r) b[(size_t)(r)*sr]
    switch( nr ) {
    case 0: break;
    case 1: v =     v+B(0); break;
    case 2: v =    (v+B(0))+ B(1); break;
    case 3: v =    (v+B(0))+(B(1)+B(2)); break;
    case 4: v =   ((v+B(0))+(B(1)+B(2)))+ B(3); break;
    case 5: v =   ((v+B(0))+(B(1)+B(2)))+(B(3)+B(4)); break;
    case 6: v =   ((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+ B(5)); break;
    case 7: v =   ((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))); break;
    case 8: v =  (((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))))+ B(7); break;
    case 9: v =  (((v+B(0))+(B(1)+B(2)))+((B(3)+B(4))+(B(5)+B(6))))+(B(7)+B(8)); break;
    }
#   undef B
    a[k] = v;
  }
}

/* F6: unload_accumulator_pipeline.cc:61-63,130-132.  Coefficients are formed
   in double (0.25 is a double constant) and rounded once to float. */
void orc_unload_accumulator_array( vpb_field_array_t * fa, const vpb_accumulator_array_t * aa ) {
  if( !fa || !aa || fa->g!=aa->g ) ORC_FAIL( "Bad args" );
  const vpb_grid_t * g = fa->g;
  const int nx = g->nx, ny = g->ny, nz = g->nz;
  const float cx = 0.25 * g->rdy * g->rdz / g->dt;
  const float cy = 0.25 * g->rdz * g->rdx / g->dt;
  const float cz = 0.25 * g->rdx * g->rdy / g->dt;
  const vpb_accumulator_t * a = aa->a;
  for( int z=1; z<=nz+1; z++ ) for( int y=1; y<=ny+1; y++ ) for( int x=1; x<=nx+1; x++ ) {
    vpb_field_t * f0 = fa->f + ORC_VOX(x,y,z);
    const vpb_accumulator_t * a0  = a + ORC_VOX(x,  y,  z  );
    const vpb_accumulator_t * ax  = a + ORC_VOX(x-1,y,  z  );
    const vpb_accumulator_t * ay  = a + ORC_VOX(x,  y-1,z  );
    const vpb_accumulator_t * az  = a + ORC_VOX(x,  y,  z-1);
    const vpb_accumulator_t * ayz = a + ORC_VOX(x,  y-1,z-1);
    const vpb_accumulator_t * azx = a + ORC_VOX(x-1,y,  z-1);
    const vpb_accumulator_t * axy = a + ORC_VOX(x-1,y-1,z  );
    f0->jfx += cx*( a0->jx[0] + ay->jx[1] + az->jx[2] + ayz->jx[3] );
    f0->jfy += cy*( a0->jy[0] + az->jy[1] + ax->jy[2] + azx->jy[3] );
    f0->jfz += cz*( a0->jz[0] + ax->jz[1] + ay->jz[2] + axy->jz[3] );
  }
}
