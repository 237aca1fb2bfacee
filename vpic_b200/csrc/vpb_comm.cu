// vpb_comm.cu -- multi-GPU transport for face exchanges, particle migration
// and reductions: one process per GPU, one NCCL communicator over NVLink 5 /
// NVSwitch.
// Reference counterpart: src/util/mp (27 MPI ports, Irecv/Issend) as used by
// src/field_advance/standard/remote.c and src/boundary/boundary_p.cc.
//
// The domain decomposition is the reference's brick-per-rank with six face
// neighbours (src/grid/partition.c:25-80); every exchange is a pair of
// messages along one axis, so the whole transport is ncclSend/ncclRecv inside
// one group on the compute stream (device buffers in, device buffers out, no
// host staging).  NCCL is bound at run time (dlopen) so the library loads on
// machines without it; a rank with a remote neighbour and no communicator
// fails loudly.
#include "vpb_runtime.h"
#include <dlfcn.h>
#include <cstring>

namespace vpb {

namespace {

// The handful of NCCL entry points we use, with their stable C signatures.
typedef struct ncclComm * ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclChar = 0, ncclInt32 = 2, ncclFloat32 = 7, ncclFloat64 = 8 };
enum { ncclSum = 0 };

struct Nccl {
  void * lib = nullptr;
  int  (*GetUniqueId)( ncclUniqueId * ) = nullptr;
  int  (*CommInitRank)( ncclComm_t *, int, ncclUniqueId, int ) = nullptr;
  int  (*CommDestroy)( ncclComm_t ) = nullptr;
  int  (*Send)( const void *, size_t, int, int, ncclComm_t, cudaStream_t ) = nullptr;
  int  (*Recv)( void *, size_t, int, int, ncclComm_t, cudaStream_t ) = nullptr;
  int  (*AllReduce)( const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t ) = nullptr;
  int  (*GroupStart)() = nullptr;
  int  (*GroupEnd)() = nullptr;
  const char * (*GetErrorString)( int ) = nullptr;
  ncclComm_t comm = nullptr;
  double * d_red = nullptr;
} N;

void load_nccl() {
  if( N.lib ) return;
  // RTLD_NOLOAD first: if the host process (e.g. torch) already carries an
  // NCCL, use that very copy.
  const char * names[] = { "libnccl.so.2", "libnccl.so" };
  for( const char * n : names ) if( !N.lib ) N.lib = dlopen( n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL );
  for( const char * n : names ) if( !N.lib ) N.lib = dlopen( n, RTLD_NOW | RTLD_GLOBAL );
  if( !N.lib ) VPB_ERROR( "multi-GPU run requested but libnccl.so.2 cannot be loaded: %s", dlerror() );
# define SYM( field, name ) *(void **)( &N.field ) = dlsym( N.lib, name ); if( !N.field ) VPB_ERROR( "libnccl has no %s", name )
  SYM( GetUniqueId, "ncclGetUniqueId" ); SYM( CommInitRank, "ncclCommInitRank" ); SYM( CommDestroy, "ncclCommDestroy" );
  SYM( Send, "ncclSend" ); SYM( Recv, "ncclRecv" ); SYM( AllReduce, "ncclAllReduce" );
  SYM( GroupStart, "ncclGroupStart" ); SYM( GroupEnd, "ncclGroupEnd" ); SYM( GetErrorString, "ncclGetErrorString" );
# undef SYM
}

#define VPB_NCCL(call) do { int e_ = (call); if( e_!=ncclSuccess ) \
  VPB_ERROR( "NCCL error %d: %s", e_, N.GetErrorString ? N.GetErrorString( e_ ) : "?" ); } while(0)

void need_comm( const char * what ) {
  if( !N.comm ) VPB_ERROR( "%s needs a remote rank (world_size=%d) but no multi-GPU transport is initialised: "
                           "call vpic_b200_comm_init on every rank first", what, g_world_size );
}

} // namespace

// Pairwise exchange along one axis.  Messages to the same peer are matched in
// issue order, so when both neighbours are the same rank (two ranks, periodic)
// the peer's "low" message must be the first thing we receive from it, and it
// lands on our HIGH side: sends go lo,hi and receives go hi,lo on every rank.
void comm_sendrecv( int rank_lo, int rank_hi,
                    const void * slo, size_t nslo, const void * shi, size_t nshi,
                    void * rlo, size_t nrlo, void * rhi, size_t nrhi ) {
  need_comm( "a face exchange" );
  Runtime & r = rt();
  VPB_NCCL( N.GroupStart() );
  if( rank_lo>=0 && nslo ) VPB_NCCL( N.Send( slo, nslo, ncclChar, rank_lo, N.comm, r.stream ) );
  if( rank_hi>=0 && nshi ) VPB_NCCL( N.Send( shi, nshi, ncclChar, rank_hi, N.comm, r.stream ) );
  if( rank_hi>=0 && nrhi ) VPB_NCCL( N.Recv( rhi, nrhi, ncclChar, rank_hi, N.comm, r.stream ) );
  if( rank_lo>=0 && nrlo ) VPB_NCCL( N.Recv( rlo, nrlo, ncclChar, rank_lo, N.comm, r.stream ) );
  VPB_NCCL( N.GroupEnd() );
  r.launches++;
}

void comm_exchange_faces( FieldMirror * m, int X, float * slo, float * shi, float * rlo, float * rhi, int nfloat ) {
  const GridMirror * gm = m->grid;
  const int rank_lo = slo ? gm->fbc[X] : -1, rank_hi = shi ? gm->fbc[X+3] : -1;
  const size_t nb = sizeof(float)*(size_t)nfloat;
  comm_sendrecv( rank_lo, rank_hi, slo, nb, shi, nb, rlo, nb, rhi, nb );
}

void comm_allsum_d( double * v, int n ) {
  need_comm( "an allreduce" );
  Runtime & r = rt();
  if( n>64 ) VPB_ERROR( "allreduce of %d doubles", n );
  if( !N.d_red ) VPB_CUDA( cudaMalloc( &N.d_red, sizeof(double)*64 ) );
  h2d( N.d_red, v, sizeof(double)*(size_t)n, r.stream );
  VPB_NCCL( N.AllReduce( N.d_red, N.d_red, (size_t)n, ncclFloat64, ncclSum, N.comm, r.stream ) );
  d2h( v, N.d_red, sizeof(double)*(size_t)n, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  r.launches++;
}

bool comm_ready() { return N.comm!=nullptr; }

void comm_shutdown() {
  if( N.comm ) { N.CommDestroy( N.comm ); N.comm = nullptr; }
  if( N.d_red ) { cudaFree( N.d_red ); N.d_red = nullptr; }
}

} // namespace vpb

using namespace vpb;

extern "C" {

// Rank 0 creates the 128-byte NCCL id; the host program ships it to the other
// ranks by whatever means it has (torch.distributed broadcast, MPI_Bcast, a
// file) and every rank then calls vpic_b200_comm_init with it.
int vpic_b200_comm_unique_id( void * out, int nbytes ) {
  if( !out || nbytes<(int)sizeof(ncclUniqueId) ) VPB_ERROR( "Bad args" );
  ensure_init(); load_nccl();
  ncclUniqueId id; VPB_NCCL( N.GetUniqueId( &id ) );
  memcpy( out, &id, sizeof(id) );
  return (int)sizeof(id);
}

void vpic_b200_comm_init( const void * id_bytes, int nbytes ) {
  if( !id_bytes || nbytes<(int)sizeof(ncclUniqueId) ) VPB_ERROR( "Bad args" );
  ensure_init(); load_nccl();
  if( N.comm ) return;
  ncclUniqueId id; memcpy( &id, id_bytes, sizeof(id) );
  VPB_NCCL( N.CommInitRank( &N.comm, g_world_size, id, g_world_rank ) );
}

int vpic_b200_comm_ready( void ) { return comm_ready() ? 1 : 0; }

} // extern "C"
