/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * Single-rank loop-back stand-in for <mpi.h>, written for this repository so
 * that the reference's sources (which include <mpi.h> from
 * src/util/mp/DMPPolicy.h:4) compile on a machine without any MPI.  It
 * implements exactly the ~20 entry points DMPPolicy.h uses.  Everything is a
 * one-rank world; MPI_Issend/MPI_Irecv/MPI_Wait form a tag-matched mailbox
 * because the reference really does send field ghost planes to itself on a
 * periodic single-rank box (src/grid/grid_comm.c:20-23).
 *
 * Header-only; all functions are static inline so every translation unit of
 * the reference that includes it gets private copies that share state through
 * one weak global table.
 */
#ifndef VPIC_B200_ORACLE_MPI_STUB_H
#define VPIC_B200_ORACLE_MPI_STUB_H

#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;

#define MPI_SUCCESS      0
#define MPI_COMM_WORLD   1
#define MPI_COMM_SELF    2
#define MPI_BYTE         1
#define MPI_CHAR         2
#define MPI_INT          3
#define MPI_LONG_LONG    4
#define MPI_DOUBLE       5
#define MPI_SUM          1

typedef struct { int count; int MPI_SOURCE; int MPI_TAG; int MPI_ERROR; } MPI_Status;
#define MPI_STATUS_IGNORE ((MPI_Status *)0)

/* A request is an index into the mailbox table (0 = null request). */
typedef int MPI_Request;

#define MPISTUB_MAX_SLOTS 256

typedef struct {
  int   in_use;      /* 1 = pending recv, 2 = pending (unmatched) send, 3 = done */
  int   is_recv;
  int   tag;
  int   size;        /* bytes posted */
  int   got;         /* bytes delivered (recv) */
  void *buf;         /* user buffer (recv) or private copy (unmatched send) */
} mpistub_slot_t;

/* One table per process; weak so that every TU agrees on it. */
__attribute__((weak)) mpistub_slot_t mpistub_slots[MPISTUB_MAX_SLOTS];

static inline int mpistub_sizeof( MPI_Datatype t ) {
  switch( t ) {
  case MPI_BYTE: case MPI_CHAR: return 1;
  case MPI_INT:                 return 4;
  case MPI_LONG_LONG:           return 8;
  case MPI_DOUBLE:              return 8;
  default:                      return 1;
  }
}

static inline int mpistub_alloc_slot( void ) {
  int i;
  for( i=1; i<MPISTUB_MAX_SLOTS; i++ ) if( !mpistub_slots[i].in_use ) return i;
  fprintf( stderr, "mpi stub: out of request slots\n" );
  abort();
  return 0;
}

static inline int MPI_Init( int *pargc, char ***pargv ) { (void)pargc; (void)pargv; return MPI_SUCCESS; }
static inline int MPI_Finalize( void ) { return MPI_SUCCESS; }
static inline int MPI_Comm_dup( MPI_Comm c, MPI_Comm *out ) { *out = c; return MPI_SUCCESS; }
static inline int MPI_Comm_free( MPI_Comm *c ) { (void)c; return MPI_SUCCESS; }
static inline int MPI_Comm_rank( MPI_Comm c, int *r ) { (void)c; *r = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size( MPI_Comm c, int *s ) { (void)c; *s = 1; return MPI_SUCCESS; }
static inline int MPI_Abort( MPI_Comm c, int reason ) { (void)c; exit( reason ? reason : 1 ); return MPI_SUCCESS; }
static inline int MPI_Barrier( MPI_Comm c ) { (void)c; return MPI_SUCCESS; }

static inline int MPI_Allreduce( const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c ) {
  (void)op; (void)c; memmove( r, s, (size_t)n*mpistub_sizeof(t) ); return MPI_SUCCESS;
}
static inline int MPI_Allgather( const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, MPI_Comm c ) {
  (void)rn; (void)rt; (void)c; memmove( r, s, (size_t)sn*mpistub_sizeof(st) ); return MPI_SUCCESS;
}
static inline int MPI_Gather( const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c ) {
  (void)rn; (void)rt; (void)root; (void)c; memmove( r, s, (size_t)sn*mpistub_sizeof(st) ); return MPI_SUCCESS;
}
static inline int MPI_Send( const void *b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c ) {
  (void)b; (void)n; (void)t; (void)dst; (void)tag; (void)c;
  fprintf( stderr, "mpi stub: blocking MPI_Send to self is not supported\n" ); abort(); return 1;
}
static inline int MPI_Recv( void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *s ) {
  (void)b; (void)n; (void)t; (void)src; (void)tag; (void)c; (void)s;
  fprintf( stderr, "mpi stub: blocking MPI_Recv from self is not supported\n" ); abort(); return 1;
}

static inline int MPI_Irecv( void *buf, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *req ) {
  int i, bytes = n*mpistub_sizeof(t), r;
  (void)src; (void)c;
  /* An unmatched send with this tag may already be waiting. */
  for( i=1; i<MPISTUB_MAX_SLOTS; i++ ) {
    mpistub_slot_t *s = &mpistub_slots[i];
    if( s->in_use==2 && s->tag==tag ) {
      int m = s->size<bytes ? s->size : bytes;
      memcpy( buf, s->buf, (size_t)m );
      free( s->buf ); s->buf = NULL; s->in_use = 3; /* send now complete */
      r = mpistub_alloc_slot();
      mpistub_slots[r].in_use = 3; mpistub_slots[r].is_recv = 1; mpistub_slots[r].tag = tag;
      mpistub_slots[r].size = bytes; mpistub_slots[r].got = m; mpistub_slots[r].buf = buf;
      *req = r;
      return MPI_SUCCESS;
    }
  }
  r = mpistub_alloc_slot();
  mpistub_slots[r].in_use = 1; mpistub_slots[r].is_recv = 1; mpistub_slots[r].tag = tag;
  mpistub_slots[r].size = bytes; mpistub_slots[r].got = 0; mpistub_slots[r].buf = buf;
  *req = r;
  return MPI_SUCCESS;
}

static inline int MPI_Issend( const void *buf, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c, MPI_Request *req ) {
  int i, bytes = n*mpistub_sizeof(t), r;
  (void)dst; (void)c;
  /* Deliver straight into a posted receive with the same tag if there is one. */
  for( i=1; i<MPISTUB_MAX_SLOTS; i++ ) {
    mpistub_slot_t *s = &mpistub_slots[i];
    if( s->in_use==1 && s->is_recv && s->tag==tag ) {
      int m = s->size<bytes ? s->size : bytes;
      memcpy( s->buf, buf, (size_t)m );
      s->got = bytes; s->in_use = 3;
      r = mpistub_alloc_slot();
      mpistub_slots[r].in_use = 3; mpistub_slots[r].is_recv = 0; mpistub_slots[r].tag = tag;
      mpistub_slots[r].size = bytes; mpistub_slots[r].buf = NULL;
      *req = r;
      return MPI_SUCCESS;
    }
  }
  r = mpistub_alloc_slot();
  mpistub_slots[r].in_use = 2; mpistub_slots[r].is_recv = 0; mpistub_slots[r].tag = tag;
  mpistub_slots[r].size = bytes; mpistub_slots[r].buf = malloc( (size_t)(bytes>0?bytes:1) );
  memcpy( mpistub_slots[r].buf, buf, (size_t)bytes );
  *req = r;
  return MPI_SUCCESS;
}

static inline int MPI_Wait( MPI_Request *req, MPI_Status *status ) {
  mpistub_slot_t *s;
  if( !req || *req<=0 || *req>=MPISTUB_MAX_SLOTS ) return MPI_SUCCESS;
  s = &mpistub_slots[*req];
  if( s->is_recv && s->in_use==1 ) {
    fprintf( stderr, "mpi stub: MPI_Wait on a receive (tag %i) that nobody sent to\n", s->tag );
    abort();
  }
  /* A send that is still unmatched (in_use==2) stays parked until its receive
     is posted; the reference always posts receives first, so this is only a
     safety net. */
  if( status ) { status->count = s->is_recv ? s->got : s->size; status->MPI_SOURCE = 0; status->MPI_TAG = s->tag; status->MPI_ERROR = 0; }
  if( s->in_use==3 ) s->in_use = 0;
  *req = 0;
  return MPI_SUCCESS;
}

static inline int MPI_Get_count( const MPI_Status *status, MPI_Datatype t, int *count ) {
  *count = status->count / mpistub_sizeof(t);
  return MPI_SUCCESS;
}

#ifdef __cplusplus
}
#endif

#endif /* VPIC_B200_ORACLE_MPI_STUB_H */
