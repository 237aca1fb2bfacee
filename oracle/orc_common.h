/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.  Shared helpers of the CPU
 * oracle (see vpic_oracle.h). */
#ifndef ORC_COMMON_H
#define ORC_COMMON_H

#include "vpic_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>

extern int orc_world_rank, orc_world_size;
extern orc_exchange_fn orc_exchange; extern void * orc_exchange_user;

#define ORC_VOX(x,y,z) VPB_VOXEL(x,y,z,nx,ny,nz)

/* The reference's DISTRIBUTE (src/util/util_base.h:191-197): split N items in
   blocks of b over P pipelines; p==P is the host's straggler share. */
static inline void orc_distribute( int N, int b, int p, int P, int * i, int * n ) {
  double t = (double)(N/b)/(double)P;
  int _i = b*(int)(t*(double)p+0.5);
  *n = (p==P) ? (N%b) : (b*(int)(t*(double)(p+1)+0.5)-_i);
  *i = _i;
}

#define ORC_FAIL(msg) do { fprintf( stderr, "oracle error %s(%d): %s\n", __FILE__, __LINE__, msg ); exit(1); } while(0)

/* True when the boundary code in g->bc[] is handled locally (reference test
   "bc<0 || bc>=world_size", e.g. src/field_advance/standard/local.c:70). */
static inline int orc_bc_is_local( int bc ) { return bc<0 || bc>=orc_world_size; }

#endif
