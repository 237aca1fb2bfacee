/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 *
 * Stand-in for <mpi.h>, written for this repository so that the reference's
 * sources (which include <mpi.h> from src/util/mp/DMPPolicy.h:4, and only
 * there: src/util/mp/mp.cc is the one translation unit that sees it) compile
 * and RUN on a machine without any MPI -- on one rank, or on N ranks as N
 * processes of one node started by oracle/mprun.py.  It implements exactly
 * the ~20 entry points DMPPolicy.h:103-343 uses.
 *
 * Transport: one shared-memory segment per job (a file under /dev/shm made by
 * the launcher; a private anonymous mapping for a lone rank):
 *
 *   header      abort flag, sense-reversing barrier
 *   staging     one 64 KB slot per rank for the collectives (every rank
 *               combines the slots in rank order, so all ranks get the same
 *               bits and a run is reproducible)
 *   rings       one single-producer / single-consumer byte ring per ordered
 *               pair (src, dst); a message is an 16-byte frame header
 *               {tag, bytes} followed by the payload, streamed through the
 *               ring in pieces, so messages may be larger than a ring
 *
 * MPI_Issend streams the whole message and returns complete (a legal, eager
 * implementation of the synchronous send as far as the reference can tell);
 * while it waits for ring space it keeps draining this rank's own incoming
 * rings, so two ranks that send each other messages larger than a ring do not
 * dead-lock.  Receives are matched by (source, tag) in posting order; messages
 * that arrive before their receive is posted are parked in a malloc'ed
 * "unexpected" list.  Self-sends (a periodic one-rank box really does send
 * ghost planes to itself, src/grid/grid_comm.c:20-23) use the ring (r, r).
 *
 * Header-only; everything is static.
 */
#ifndef VPIC_B200_ORACLE_MPI_STUB_H
#define VPIC_B200_ORACLE_MPI_STUB_H

#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <stdint.h>
#include <sched.h>
#include <fcntl.h>
#include <unistd.h>
#include <sys/mman.h>

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

/* A request is an index into the request table (0 = null request). */
typedef int MPI_Request;

/* ---- layout of the shared segment (oracle/mprun.py computes the same size) ---- */
#define MPISTUB_HDR_BYTES   4096
#define MPISTUB_COLL_BYTES  65536                   /* staging slot per rank */
#define MPISTUB_RING_BYTES  (4u<<20)                /* payload bytes per ring (power of two) */
#define MPISTUB_RING_HDR    128                     /* head and tail on their own cache lines */
#define MPISTUB_MAX_REQ     256

typedef struct { volatile int abort_flag; volatile int bar_count; volatile int bar_sense; } mpistub_hdr_t;
typedef struct { volatile uint64_t head; char pad0[56]; volatile uint64_t tail; char pad1[56]; } mpistub_ring_t;

typedef struct mpistub_unexp {           /* a message that arrived before its receive was posted */
  struct mpistub_unexp * next;
  int src, tag, size, got, complete, claimed;
  char * data;
} mpistub_unexp_t;

typedef struct {
  int   state;       /* 0 free, 1 pending receive, 2 complete */
  int   is_recv, src, tag;
  int   size;        /* bytes posted */
  int   got;         /* bytes of the matched message */
  char *buf;
  mpistub_unexp_t * from;   /* bound to a parked message that is still arriving */
} mpistub_req_t;

typedef struct {     /* progress of the message at the front of one incoming ring */
  int in_msg, tag, size, got;
  char * dst; int dst_cap;
  mpistub_req_t * req; mpistub_unexp_t * unexp;
} mpistub_in_t;

static struct {
  int rank, size, local_sense;
  char * base; size_t bytes;
  mpistub_req_t req[MPISTUB_MAX_REQ];
  mpistub_unexp_t * unexp_head, * unexp_tail;
  mpistub_in_t * in;          /* [size] */
} mpistub;

static inline size_t mpistub_segment_bytes( int size ) {
  return (size_t)MPISTUB_HDR_BYTES + (size_t)size*MPISTUB_COLL_BYTES
       + (size_t)size*size*( (size_t)MPISTUB_RING_HDR + MPISTUB_RING_BYTES );
}
static inline mpistub_hdr_t * mpistub_hdr( void ) { return (mpistub_hdr_t *)mpistub.base; }
static inline char * mpistub_slot( int r ) { return mpistub.base + MPISTUB_HDR_BYTES + (size_t)r*MPISTUB_COLL_BYTES; }
static inline mpistub_ring_t * mpistub_ring( int src, int dst ) {
  return (mpistub_ring_t *)( mpistub.base + MPISTUB_HDR_BYTES + (size_t)mpistub.size*MPISTUB_COLL_BYTES
                             + ( (size_t)src*mpistub.size + dst )*( (size_t)MPISTUB_RING_HDR + MPISTUB_RING_BYTES ) );
}
static inline char * mpistub_ring_data( mpistub_ring_t * r ) { return (char *)r + MPISTUB_RING_HDR; }

static inline void mpistub_die( const char * what ) {
  fprintf( stderr, "mpi stub[%d]: %s\n", mpistub.rank, what ); fflush( stderr );
  if( mpistub.base ) __atomic_store_n( &mpistub_hdr()->abort_flag, 1, __ATOMIC_RELEASE );
  _exit( 1 );
}
static inline void mpistub_relax( void ) {
  if( __atomic_load_n( &mpistub_hdr()->abort_flag, __ATOMIC_ACQUIRE ) ) _exit( 1 );   /* another rank aborted */
  sched_yield();
}

static inline int mpistub_sizeof( MPI_Datatype t ) {
  switch( t ) {
  case MPI_BYTE: case MPI_CHAR: return 1;
  case MPI_INT:                 return 4;
  case MPI_LONG_LONG:           return 8;
  case MPI_DOUBLE:              return 8;
  default:                      return 1;
  }
}

/* copy n bytes out of / into a ring at absolute position pos (wraps) */
static inline void mpistub_ring_get( mpistub_ring_t * r, uint64_t pos, void * dst, size_t n ) {
  const size_t o = (size_t)( pos & ( MPISTUB_RING_BYTES-1 ) ), first = n<MPISTUB_RING_BYTES-o ? n : MPISTUB_RING_BYTES-o;
  memcpy( dst, mpistub_ring_data( r )+o, first );
  if( n>first ) memcpy( (char *)dst+first, mpistub_ring_data( r ), n-first );
}
static inline void mpistub_ring_put( mpistub_ring_t * r, uint64_t pos, const void * src, size_t n ) {
  const size_t o = (size_t)( pos & ( MPISTUB_RING_BYTES-1 ) ), first = n<MPISTUB_RING_BYTES-o ? n : MPISTUB_RING_BYTES-o;
  memcpy( mpistub_ring_data( r )+o, src, first );
  if( n>first ) memcpy( mpistub_ring_data( r ), (const char *)src+first, n-first );
}

/* Drain every incoming ring as far as it goes. */
static inline void mpistub_progress( void ) {
  int s;
  for( s=0; s<mpistub.size; s++ ) {
    mpistub_ring_t * r = mpistub_ring( s, mpistub.rank );
    mpistub_in_t * in = &mpistub.in[s];
    for(;;) {
      const uint64_t head = __atomic_load_n( &r->head, __ATOMIC_ACQUIRE );
      uint64_t tail = r->tail;
      uint64_t avail = head - tail;
      if( !in->in_msg ) {
        int64_t frame[2]; int k;
        if( avail<16 ) break;
        mpistub_ring_get( r, tail, frame, 16 );
        tail += 16; avail -= 16;
        in->in_msg = 1; in->tag = (int)frame[0]; in->size = (int)frame[1]; in->got = 0; in->req = NULL; in->unexp = NULL;
        for( k=1; k<MPISTUB_MAX_REQ; k++ ) {            /* oldest posted receive for (s, tag) */
          mpistub_req_t * q = &mpistub.req[k];
          if( q->state==1 && q->is_recv && !q->from && q->src==s && q->tag==in->tag && q->got<0 ) { in->req = q; break; }
        }
        if( in->req ) {
          /* requests are scanned by index, not by posting time; the reference never has two receives with the
             same (source, tag) outstanding (one port per direction), so the first hit is the only one */
          in->req->got = in->size; in->dst = in->req->buf; in->dst_cap = in->req->size;
        } else {
          mpistub_unexp_t * u = (mpistub_unexp_t *)calloc( 1, sizeof(*u) );
          u->src = s; u->tag = in->tag; u->size = in->size; u->data = (char *)malloc( in->size>0 ? (size_t)in->size : 1 );
          if( mpistub.unexp_tail ) mpistub.unexp_tail->next = u; else mpistub.unexp_head = u;
          mpistub.unexp_tail = u;
          in->unexp = u; in->dst = u->data; in->dst_cap = in->size;
        }
        __atomic_store_n( &r->tail, tail, __ATOMIC_RELEASE );
      }
      {
        const int left = in->size - in->got;
        /* the payload is padded to 8 bytes in the ring */
        int take = left<(int64_t)avail ? left : (int)avail;
        if( take>0 ) {
          const int room = in->dst_cap - in->got;
          const int keep = take<room ? take : ( room>0 ? room : 0 );
          if( keep>0 ) mpistub_ring_get( r, tail, in->dst + in->got, (size_t)keep );
          in->got += take; tail += (uint64_t)take;
          if( in->unexp ) in->unexp->got = in->got;
        }
        if( in->got==in->size ) {
          const uint64_t pad = (uint64_t)( ( 8 - ( in->size & 7 ) ) & 7 );
          if( head - tail < pad ) { __atomic_store_n( &r->tail, tail, __ATOMIC_RELEASE ); break; }   /* padding not written yet */
          tail += pad;
          if( in->req ) in->req->state = 2;
          if( in->unexp ) in->unexp->complete = 1;
          in->in_msg = 0;
          __atomic_store_n( &r->tail, tail, __ATOMIC_RELEASE );
          continue;
        }
        __atomic_store_n( &r->tail, tail, __ATOMIC_RELEASE );
        if( take<=0 ) break;
      }
    }
  }
  /* receives bound to parked messages that have finished arriving */
  {
    int k;
    for( k=1; k<MPISTUB_MAX_REQ; k++ ) {
      mpistub_req_t * q = &mpistub.req[k];
      if( q->state==1 && q->from && q->from->complete ) {
        mpistub_unexp_t * u = q->from, ** pp;
        memcpy( q->buf, u->data, (size_t)( u->size<q->size ? u->size : q->size ) );
        q->got = u->size; q->state = 2; q->from = NULL;
        for( pp=&mpistub.unexp_head; *pp; pp=&(*pp)->next ) if( *pp==u ) { *pp = u->next; break; }
        mpistub.unexp_tail = NULL;
        { mpistub_unexp_t * t; for( t=mpistub.unexp_head; t; t=t->next ) mpistub.unexp_tail = t; }
        free( u->data ); free( u );
      }
    }
  }
}

static inline void mpistub_barrier( void ) {
  mpistub_hdr_t * h = mpistub_hdr();
  const int s = !mpistub.local_sense;
  mpistub.local_sense = s;
  if( __atomic_add_fetch( &h->bar_count, 1, __ATOMIC_ACQ_REL )==mpistub.size ) {
    __atomic_store_n( &h->bar_count, 0, __ATOMIC_RELAXED );
    __atomic_store_n( &h->bar_sense, s, __ATOMIC_RELEASE );
  } else while( __atomic_load_n( &h->bar_sense, __ATOMIC_ACQUIRE )!=s ) { mpistub_progress(); mpistub_relax(); }
}

static inline int mpistub_alloc_req( void ) {
  int i;
  for( i=1; i<MPISTUB_MAX_REQ; i++ ) if( !mpistub.req[i].state ) return i;
  mpistub_die( "out of request slots" );
  return 0;
}

static inline int MPI_Init( int *pargc, char ***pargv ) {
  const char * er = getenv( "VPB_MPI_RANK" ), * es = getenv( "VPB_MPI_SIZE" ), * ef = getenv( "VPB_MPI_SHM" );
  (void)pargc; (void)pargv;
  mpistub.rank = 0; mpistub.size = 1;
  if( er && es && ef && atoi( es )>1 ) {
    int fd;
    mpistub.rank = atoi( er ); mpistub.size = atoi( es );
    mpistub.bytes = mpistub_segment_bytes( mpistub.size );
    fd = open( ef, O_RDWR );
    if( fd<0 ) { perror( "mpi stub: open VPB_MPI_SHM" ); _exit( 1 ); }
    mpistub.base = (char *)mmap( NULL, mpistub.bytes, PROT_READ|PROT_WRITE, MAP_SHARED, fd, 0 );
    close( fd );
  } else {
    mpistub.bytes = mpistub_segment_bytes( 1 );
    mpistub.base = (char *)mmap( NULL, mpistub.bytes, PROT_READ|PROT_WRITE, MAP_PRIVATE|MAP_ANONYMOUS, -1, 0 );
  }
  if( mpistub.base==(char *)MAP_FAILED ) { perror( "mpi stub: mmap" ); _exit( 1 ); }
  mpistub.in = (mpistub_in_t *)calloc( (size_t)mpistub.size, sizeof(mpistub_in_t) );
  mpistub.local_sense = 0;
  if( mpistub.size>1 ) mpistub_barrier();
  return MPI_SUCCESS;
}
static inline int MPI_Finalize( void ) { if( mpistub.base && mpistub.size>1 ) mpistub_barrier(); return MPI_SUCCESS; }
static inline int MPI_Comm_dup( MPI_Comm c, MPI_Comm *out ) { *out = c; return MPI_SUCCESS; }
static inline int MPI_Comm_free( MPI_Comm *c ) { (void)c; return MPI_SUCCESS; }
static inline int MPI_Comm_rank( MPI_Comm c, int *r ) { *r = c==MPI_COMM_SELF ? 0 : mpistub.rank; return MPI_SUCCESS; }
static inline int MPI_Comm_size( MPI_Comm c, int *s ) { *s = c==MPI_COMM_SELF ? 1 : mpistub.size; return MPI_SUCCESS; }
static inline int MPI_Abort( MPI_Comm c, int reason ) {
  (void)c;
  if( mpistub.base ) __atomic_store_n( &mpistub_hdr()->abort_flag, 1, __ATOMIC_RELEASE );
  exit( reason ? reason : 1 );
  return MPI_SUCCESS;
}
static inline int MPI_Barrier( MPI_Comm c ) { (void)c; if( mpistub.size>1 ) mpistub_barrier(); return MPI_SUCCESS; }

/* Collectives: every rank writes its contribution to its staging slot; after a
   barrier every rank reads all slots in rank order; a second barrier protects
   the slots from the next collective. */
static inline void mpistub_stage( const void * s, size_t bytes ) {
  if( bytes>MPISTUB_COLL_BYTES ) mpistub_die( "collective larger than the staging slot" );
  memcpy( mpistub_slot( mpistub.rank ), s, bytes );
  mpistub_barrier();
}
static inline int MPI_Allreduce( const void *s, void *r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c ) {
  int k, i;
  (void)op; (void)c;
  if( mpistub.size==1 ) { memmove( r, s, (size_t)n*mpistub_sizeof(t) ); return MPI_SUCCESS; }
  mpistub_stage( s, (size_t)n*mpistub_sizeof(t) );
  if( t==MPI_DOUBLE ) {
    double * out = (double *)r;
    for( i=0; i<n; i++ ) { double a = 0; for( k=0; k<mpistub.size; k++ ) a += ((const double *)mpistub_slot( k ))[i]; out[i] = a; }
  } else if( t==MPI_INT ) {
    int * out = (int *)r;
    for( i=0; i<n; i++ ) { int a = 0; for( k=0; k<mpistub.size; k++ ) a += ((const int *)mpistub_slot( k ))[i]; out[i] = a; }
  } else if( t==MPI_LONG_LONG ) {
    long long * out = (long long *)r;
    for( i=0; i<n; i++ ) { long long a = 0; for( k=0; k<mpistub.size; k++ ) a += ((const long long *)mpistub_slot( k ))[i]; out[i] = a; }
  } else mpistub_die( "MPI_Allreduce: unsupported datatype" );
  mpistub_barrier();
  return MPI_SUCCESS;
}
static inline int MPI_Allgather( const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, MPI_Comm c ) {
  const size_t b = (size_t)sn*mpistub_sizeof(st);
  int k;
  (void)rn; (void)rt; (void)c;
  if( mpistub.size==1 ) { memmove( r, s, b ); return MPI_SUCCESS; }
  mpistub_stage( s, b );
  for( k=0; k<mpistub.size; k++ ) memcpy( (char *)r + (size_t)k*b, mpistub_slot( k ), b );
  mpistub_barrier();
  return MPI_SUCCESS;
}
static inline int MPI_Gather( const void *s, int sn, MPI_Datatype st, void *r, int rn, MPI_Datatype rt, int root, MPI_Comm c ) {
  const size_t b = (size_t)sn*mpistub_sizeof(st);
  int k;
  (void)rn; (void)rt; (void)c;
  if( mpistub.size==1 ) { memmove( r, s, b ); return MPI_SUCCESS; }
  mpistub_stage( s, b );
  if( mpistub.rank==root ) for( k=0; k<mpistub.size; k++ ) memcpy( (char *)r + (size_t)k*b, mpistub_slot( k ), b );
  mpistub_barrier();
  return MPI_SUCCESS;
}

static inline int MPI_Irecv( void *buf, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *req ) {
  const int bytes = n*mpistub_sizeof(t), r = mpistub_alloc_req();
  mpistub_req_t * q = &mpistub.req[r];
  mpistub_unexp_t * u;
  (void)c;
  if( src<0 || src>=mpistub.size ) mpistub_die( "MPI_Irecv: bad source" );
  q->state = 1; q->is_recv = 1; q->src = src; q->tag = tag; q->size = bytes; q->got = -1; q->buf = (char *)buf; q->from = NULL;
  /* the oldest parked message from (src, tag), complete or still arriving */
  for( u=mpistub.unexp_head; u; u=u->next ) if( u->src==src && u->tag==tag && !u->claimed ) { u->claimed = 1; q->from = u; q->got = u->size; break; }
  mpistub_progress();
  *req = r;
  return MPI_SUCCESS;
}

static inline int MPI_Issend( const void *buf, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c, MPI_Request *req ) {
  const int bytes = n*mpistub_sizeof(t);
  const int padded = ( bytes + 7 ) & ~7;
  mpistub_ring_t * r;
  int64_t frame[2];
  int sent = 0, hdr_done = 0, k;
  static const char zeros[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
  (void)c;
  if( dst<0 || dst>=mpistub.size ) mpistub_die( "MPI_Issend: bad destination" );
  r = mpistub_ring( mpistub.rank, dst );
  frame[0] = tag; frame[1] = bytes;
  while( !hdr_done || sent<padded ) {
    uint64_t head = r->head;
    const uint64_t tail = __atomic_load_n( &r->tail, __ATOMIC_ACQUIRE );
    uint64_t space = MPISTUB_RING_BYTES - ( head - tail );
    int moved = 0;
    if( !hdr_done && space>=16 ) { mpistub_ring_put( r, head, frame, 16 ); head += 16; space -= 16; hdr_done = 1; moved = 1; }
    if( hdr_done && sent<padded && space>0 ) {
      int take = padded-sent<(int64_t)space ? padded-sent : (int)space;
      const int real = sent<bytes ? ( take<bytes-sent ? take : bytes-sent ) : 0;
      if( real>0 ) mpistub_ring_put( r, head, (const char *)buf+sent, (size_t)real );
      if( take>real ) mpistub_ring_put( r, head+(uint64_t)real, zeros, (size_t)( take-real ) );
      head += (uint64_t)take; sent += take; moved = 1;
    }
    if( moved ) __atomic_store_n( &r->head, head, __ATOMIC_RELEASE );
    else { mpistub_progress(); mpistub_relax(); }
  }
  k = mpistub_alloc_req();
  mpistub.req[k].state = 2; mpistub.req[k].is_recv = 0; mpistub.req[k].src = dst; mpistub.req[k].tag = tag;
  mpistub.req[k].size = bytes; mpistub.req[k].got = bytes; mpistub.req[k].buf = NULL; mpistub.req[k].from = NULL;
  *req = k;
  return MPI_SUCCESS;
}

static inline int MPI_Wait( MPI_Request *req, MPI_Status *status ) {
  mpistub_req_t * q;
  if( !req || *req<=0 || *req>=MPISTUB_MAX_REQ ) return MPI_SUCCESS;
  q = &mpistub.req[*req];
  while( q->state==1 ) { mpistub_progress(); if( q->state==1 ) mpistub_relax(); }
  if( q->is_recv && q->got>q->size ) mpistub_die( "message longer than the posted receive" );
  if( status ) { status->count = q->got; status->MPI_SOURCE = q->src; status->MPI_TAG = q->tag; status->MPI_ERROR = 0; }
  q->state = 0;
  *req = 0;
  return MPI_SUCCESS;
}

static inline int MPI_Send( const void *b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c ) {
  MPI_Request r; MPI_Issend( b, n, t, dst, tag, c, &r ); return MPI_Wait( &r, MPI_STATUS_IGNORE );
}
static inline int MPI_Recv( void *b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *s ) {
  MPI_Request r; MPI_Irecv( b, n, t, src, tag, c, &r ); return MPI_Wait( &r, s );
}

static inline int MPI_Get_count( const MPI_Status *status, MPI_Datatype t, int *count ) {
  *count = status->count / mpistub_sizeof(t);
  return MPI_SUCCESS;
}

#ifdef __cplusplus
}
#endif

#endif /* VPIC_B200_ORACLE_MPI_STUB_H */
