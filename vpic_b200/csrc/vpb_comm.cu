// vpb_comm.cu -- multi-GPU transport for face exchanges, particle migration
// and reductions: one process per GPU, one NCCL communicator over NVLink 5 /
// NVSwitch.
// Reference counterpart: src/util/mp (27 MPI ports, Irecv/Issend) as used by
// src/field_advance/standard/remote.c and src/boundary/boundary_p.cc.
//
// The domain decomposition is the reference's brick-per-rank with six face
// neighbours (src/grid/partition.c:25-80); every exchange is a pair of
// messages along one axis.  NCCL is bound at run time (dlopen) so the library
// loads on machines without it; a rank with a remote neighbour and no
// communicator fails loudly.
//
// Data path ("put" transport, the default when every rank's mailbox can be
// mapped): the messages of the step are sub-MB and latency-bound, so they do
// not go through ncclSend/ncclRecv.  Every rank owns a MAILBOX in device
// memory, exported with cudaIpc and mapped by its neighbours over NVLink peer
// memory: slot (source rank, direction, parity).  One kernel stores the
// outgoing messages straight into the neighbours' mailboxes (128-bit stores
// over NVLink), fences, and the last block publishes the message's sequence
// number in the mailbox's flag word (st.release.sys); one kernel on the
// receiving rank acquires the flag and copies the payload out.  No host
// round-trip, no proxy thread, two launches per exchange.  Every exchange is
// bidirectional per neighbour pair (zero-byte messages still publish their
// flag), which is what makes two parity slots enough: a rank cannot send
// message k+2 before it has received k+1, which its neighbour sent after
// consuming k.  The field exchanges go one step further (comm_link_reserve):
// the plane kernels of vpb_fields.cu pack straight into the neighbour's slot
// and publish, and unpack straight out of their own slot after acquiring the
// flag -- no k_put / k_get and no staging buffer on that path.  NCCL remains
// for the bootstrap (all-gather of the IPC
// handles), the energy all-reduce and messages larger than a mailbox slot.
// VPIC_B200_TRANSPORT=nccl selects the ncclSend/ncclRecv path for A/B runs.
#include "vpb_runtime.h"
#include <dlfcn.h>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <algorithm>

namespace vpb {

void bootstrap_from_host_program();

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
  int  (*AllGather)( const void *, void *, size_t, int, ncclComm_t, cudaStream_t ) = nullptr;
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
  SYM( Send, "ncclSend" ); SYM( Recv, "ncclRecv" ); SYM( AllReduce, "ncclAllReduce" ); SYM( AllGather, "ncclAllGather" );
  SYM( GroupStart, "ncclGroupStart" ); SYM( GroupEnd, "ncclGroupEnd" ); SYM( GetErrorString, "ncclGetErrorString" );
# undef SYM
}

#define VPB_NCCL(call) do { int e_ = (call); if( e_!=ncclSuccess ) \
  VPB_ERROR( "NCCL error %d: %s", e_, N.GetErrorString ? N.GetErrorString( e_ ) : "?" ); } while(0)

bool mailbox_on();

void need_comm( const char * what ) {
  if( !N.comm && !mailbox_on() ) bootstrap_from_host_program();
  if( !N.comm && !mailbox_on() )
    VPB_ERROR( "%s needs a remote rank (world_size=%d) but no multi-GPU transport is initialised: "
               "call vpic_b200_comm_init (or vpic_b200_comm_init_handles) on every rank first", what, g_world_size );
}

// ---- put transport over NVLink peer memory -------------------------------------
constexpr size_t MB_SLOT  = 8u<<20;            // payload bytes per mailbox slot
constexpr int    MB_FLAGW = 32;                // one 128 B line per flag

struct Mailbox {
  bool     on = false, tried = false;
  char *   mine = nullptr;                     // this rank's mailbox (cudaMalloc, IPC-exported)
  std::vector<char *> peer;                    // mapped mailboxes of the other ranks (nullptr until first use)
  std::vector<cudaIpcMemHandle_t> handle;
  std::vector<unsigned> sent, rcvd;            // per (rank, direction): messages put to / taken from that slot pair
  unsigned * d_done = nullptr;                 // block counters of the put kernel (2)
  unsigned * d_done_many = nullptr;            // block counters for fused pack kernels (one per batched operation)
  int *    d_err = nullptr;                    // set by a receive that timed out
  size_t   bytes = 0;
  cudaIpcMemHandle_t my_handle; bool exported = false;
  char **  d_peers = nullptr;                  // device copy of every rank's mapped mailbox address (all-reduce)
  double * d_red = nullptr;                    // all-reduce operand / result
  unsigned red_seq = 0;
} MB;

bool mailbox_on() { return MB.on; }

// mailbox layout of one rank: [world][2 directions][2 parities] payload slots, then the flags
__host__ __device__ inline size_t mb_slot_off( int src, int dir, int par ) { return ( ( (size_t)src*2 + dir )*2 + par )*MB_SLOT; }
inline size_t mb_flag_off( int world, int src, int dir, int par ) {
  return (size_t)world*4*MB_SLOT + ( ( (size_t)src*2 + dir )*2 + par )*MB_FLAGW*sizeof(unsigned);
}

// ... then the all-reduce area: per (source rank, parity) RED_MAX doubles followed by a flag line
constexpr int RED_MAX = 64;
constexpr size_t RED_SLOT = RED_MAX*sizeof(double) + 128;
__host__ __device__ inline size_t mb_red_off( int world, int src, int par ) {
  return (size_t)world*4*MB_SLOT + (size_t)world*4*MB_FLAGW*sizeof(unsigned) + ( (size_t)src*2 + par )*RED_SLOT;
}

struct PutMsg { const char * src; char * dst; unsigned * flag; unsigned nbytes, seq; };

// All-reduce (sum) of n <= RED_MAX doubles over the mailboxes: every rank stores
// its vector into its own slot of EVERY rank's mailbox (peer stores over NVLink)
// and publishes; then every rank adds the world's vectors out of its own mailbox
// IN RANK ORDER, so all ranks hold the same bits and a run is reproducible (the
// order MPI_Allreduce uses is implementation-defined; rank order is what the
// test stand-in for MPI does).
__global__ void k_allsum_put( char * const * __restrict__ peers, int world, int me, int par, unsigned seq,
                              const double * __restrict__ v, int n ) {
  for( int t=threadIdx.x; t<world*n; t+=blockDim.x ) {
    const int p = t/n, i = t - p*n;
    reinterpret_cast<double *>( peers[p] + mb_red_off( world, me, par ) )[i] = v[i];
  }
  __threadfence_system();
  __syncthreads();
  for( int p=threadIdx.x; p<world; p+=blockDim.x )
    asm volatile( "st.release.sys.global.u32 [%0], %1;" :: "l"( peers[p] + mb_red_off( world, me, par ) + RED_MAX*sizeof(double) ), "r"(seq) : "memory" );
}

__global__ void k_allsum_get( const char * __restrict__ mine, int world, int par, unsigned seq, double * __restrict__ out, int n,
                              int * __restrict__ err, long long timeout_clocks ) {
  if( threadIdx.x==0 ) {
    const long long t0 = clock64();
    for( int s=0; s<world; s++ ) {
      const char * flag = mine + mb_red_off( world, s, par ) + RED_MAX*sizeof(double);
      unsigned v;
      do {
        asm volatile( "ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory" );
        if( (int)( v-seq )>=0 ) break;
        if( clock64()-t0>timeout_clocks ) {
          printf( "[vpic_b200] put transport: all-reduce %u never heard from rank %d (flag %u)\n", seq, s, v );
          atomicExch( err, 1 ); __trap();
        }
        __nanosleep( 64 );
      } while( true );
    }
  }
  __syncthreads();
  if( (int)threadIdx.x<n ) {
    double a = 0;
    for( int s=0; s<world; s++ ) a += reinterpret_cast<const double *>( mine + mb_red_off( world, s, par ) )[threadIdx.x];
    out[threadIdx.x] = a;
  }
}

// blockIdx.y picks the message.  Payload first, then a system-scope fence per
// thread; the last block to finish publishes the sequence number.
__global__ void k_put( PutMsg m0, PutMsg m1, unsigned * __restrict__ done ) {
  const PutMsg m = blockIdx.y ? m1 : m0;
  if( !m.flag ) return;
  const unsigned n16 = m.nbytes>>4;
  if( ( ( (size_t)m.src | (size_t)m.dst ) & 15 )==0 ) {
    const uint4 * s = reinterpret_cast<const uint4 *>( m.src ); uint4 * d = reinterpret_cast<uint4 *>( m.dst );
    for( unsigned i=blockIdx.x*blockDim.x+threadIdx.x; i<n16; i+=gridDim.x*blockDim.x ) d[i] = s[i];
    for( unsigned i=( n16<<4 )+blockIdx.x*blockDim.x+threadIdx.x; i<m.nbytes; i+=gridDim.x*blockDim.x ) m.dst[i] = m.src[i];
  } else
    for( unsigned i=blockIdx.x*blockDim.x+threadIdx.x; i<m.nbytes; i+=gridDim.x*blockDim.x ) m.dst[i] = m.src[i];
  __threadfence_system();
  __syncthreads();
  if( threadIdx.x==0 ) {
    if( atomicAdd( done+blockIdx.y, 1u )==gridDim.x-1 ) {
      done[blockIdx.y] = 0;
      asm volatile( "st.release.sys.global.u32 [%0], %1;" :: "l"(m.flag), "r"(m.seq) : "memory" );
    }
  }
}

struct GetMsg { char * dst; const char * src; const unsigned * flag; unsigned nbytes, seq; };

// ~10 min by default: the neighbour is gone.  VPIC_B200_PUT_TIMEOUT_S shortens it (tests).
long long put_timeout_clocks() {
  static long long c = 0;
  if( !c ) { const char * e = getenv( "VPIC_B200_PUT_TIMEOUT_S" ); const double sec = e ? atof( e ) : 600.0; c = (long long)( ( sec>0 ? sec : 600.0 )*2.0e9 ); }
  return c;
}

__global__ void k_get( GetMsg m0, GetMsg m1, int * __restrict__ err, long long timeout_clocks ) {
  const GetMsg m = blockIdx.y ? m1 : m0;
  if( !m.flag ) return;
  if( threadIdx.x==0 ) {
    unsigned v; const long long t0 = clock64();
    do {
      asm volatile( "ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(m.flag) : "memory" );
      if( (int)( v-m.seq )>=0 ) break;
      if( clock64()-t0>timeout_clocks ) {
        printf( "[vpic_b200] put transport: message %u never arrived (flag %u)\n", m.seq, v );
        atomicExch( err, 1 ); __trap();
      }
      __nanosleep( 64 );
    } while( true );
  }
  __syncthreads();
  const unsigned n16 = m.nbytes>>4;
  if( ( ( (size_t)m.src | (size_t)m.dst ) & 15 )==0 ) {
    const uint4 * s = reinterpret_cast<const uint4 *>( m.src ); uint4 * d = reinterpret_cast<uint4 *>( m.dst );
    for( unsigned i=blockIdx.x*blockDim.x+threadIdx.x; i<n16; i+=gridDim.x*blockDim.x ) d[i] = s[i];
    for( unsigned i=( n16<<4 )+blockIdx.x*blockDim.x+threadIdx.x; i<m.nbytes; i+=gridDim.x*blockDim.x ) m.dst[i] = m.src[i];
  } else
    for( unsigned i=blockIdx.x*blockDim.x+threadIdx.x; i<m.nbytes; i+=gridDim.x*blockDim.x ) m.dst[i] = m.src[i];
}

// Allocate this rank's mailbox and export it (cudaIpc).  Returns whether the export worked.
bool mailbox_alloc() {
  if( MB.mine ) return MB.exported;
  const int W = g_world_size;
  MB.bytes = (size_t)W*4*MB_SLOT + (size_t)W*4*MB_FLAGW*sizeof(unsigned) + (size_t)W*2*RED_SLOT;
  VPB_CUDA( cudaMalloc( &MB.mine, MB.bytes ) );
  VPB_CUDA( cudaMemset( MB.mine + (size_t)W*4*MB_SLOT, 0, MB.bytes - (size_t)W*4*MB_SLOT ) );
  VPB_CUDA( cudaMalloc( &MB.d_done, sizeof(unsigned)*2 ) ); VPB_CUDA( cudaMemset( MB.d_done, 0, sizeof(unsigned)*2 ) );
  VPB_CUDA( cudaMalloc( &MB.d_err, sizeof(int) ) );         VPB_CUDA( cudaMemset( MB.d_err, 0, sizeof(int) ) );
  VPB_CUDA( cudaMalloc( &MB.d_red, sizeof(double)*RED_MAX ) );
  memset( &MB.my_handle, 0, sizeof(MB.my_handle) );
  MB.exported = cudaIpcGetMemHandle( &MB.my_handle, MB.mine )==cudaSuccess;
  if( !MB.exported ) cudaGetLastError();
  return MB.exported;
}

// Every rank's { ok, handle } record is known: switch the put transport on if all exports worked.
constexpr size_t MB_REC = 16 + sizeof(cudaIpcMemHandle_t);
void mailbox_install( const char * recs ) {
  const int W = g_world_size;
  MB.handle.resize( W ); MB.peer.assign( W, nullptr ); MB.sent.assign( 2*W, 0 ); MB.rcvd.assign( 2*W, 0 );
  bool all = true;
  for( int k=0; k<W; k++ ) { int f; memcpy( &f, recs + MB_REC*k, sizeof(int) ); all = all && f; memcpy( &MB.handle[k], recs + MB_REC*k + 16, sizeof(cudaIpcMemHandle_t) ); }
  MB.on = all;                                 // every rank reaches the same verdict
  if( !all && g_world_rank==0 ) VPB_WARNING( "cudaIpc export of the mailbox failed on some rank: face exchanges use ncclSend/ncclRecv" );
}

// Allocate and export this rank's mailbox, gather every rank's handle through NCCL.  Collective.
void mailbox_setup() {
  if( MB.tried ) return;
  MB.tried = true;
  if( const char * e = getenv( "VPIC_B200_TRANSPORT" ) ) if( !strcmp( e, "nccl" ) ) return;
  Runtime & r = rt();
  const int W = g_world_size;
  const int ok = mailbox_alloc() ? 1 : 0;
  std::vector<char> host( MB_REC*W, 0 );
  memcpy( &host[MB_REC*g_world_rank], &ok, sizeof(int) ); memcpy( &host[MB_REC*g_world_rank+16], &MB.my_handle, sizeof(MB.my_handle) );
  char * d_all = nullptr;
  VPB_CUDA( cudaMalloc( &d_all, MB_REC*W ) );
  VPB_CUDA( cudaMemcpyAsync( d_all, host.data(), MB_REC*W, cudaMemcpyHostToDevice, r.stream ) );
  VPB_NCCL( N.AllGather( d_all + MB_REC*g_world_rank, d_all, MB_REC, ncclChar, N.comm, r.stream ) );
  VPB_CUDA( cudaMemcpyAsync( host.data(), d_all, MB_REC*W, cudaMemcpyDeviceToHost, r.stream ) );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  VPB_CUDA( cudaFree( d_all ) );
  mailbox_install( host.data() );
}

char * mailbox_of( int rank ) {
  if( rank==g_world_rank ) return MB.mine;
  if( !MB.peer[rank] ) {
    void * q = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle( &q, MB.handle[rank], cudaIpcMemLazyEnablePeerAccess );
    if( e!=cudaSuccess ) VPB_ERROR( "cannot map the mailbox of rank %d over NVLink peer memory (%s); set VPIC_B200_TRANSPORT=nccl", rank, cudaGetErrorString( e ) );
    MB.peer[rank] = (char *)q;
  }
  return MB.peer[rank];
}

bool put_sendrecv( int rank_lo, int rank_hi, const void * slo, size_t nslo, const void * shi, size_t nshi,
                   void * rlo, size_t nrlo, void * rhi, size_t nrhi ) {
  if( !MB.on ) return false;
  if( rank_lo<0 && rank_hi<0 ) return true;
  Runtime & r = rt();
  const int W = g_world_size, me = g_world_rank;
  PutMsg p[2] = {}; GetMsg q[2] = {};
  // my "lo" message lands in rank_lo's slot (me, direction 0); my "hi" message in rank_hi's slot (me, 1)
  const int to[2] = { rank_lo, rank_hi }; const void * sb[2] = { slo, shi }; const size_t sn[2] = { nslo, nshi };
  for( int d=0; d<2; d++ ) if( to[d]>=0 ) {
    const unsigned k = MB.sent[2*to[d]+d]++;
    char * box = mailbox_of( to[d] );
    p[d].src = (const char *)sb[d]; p[d].dst = box + mb_slot_off( me, d, k&1 ); p[d].nbytes = (unsigned)sn[d]; p[d].seq = k+1;
    p[d].flag = reinterpret_cast<unsigned *>( box + mb_flag_off( W, me, d, k&1 ) );
  }
  // what arrives from rank_lo was its "hi" message (direction 1); from rank_hi its "lo" message (direction 0)
  const int from[2] = { rank_lo, rank_hi }; void * rb[2] = { rlo, rhi }; const size_t rn[2] = { nrlo, nrhi };
  for( int d=0; d<2; d++ ) if( from[d]>=0 ) {
    const int dir = 1-d;
    const unsigned k = MB.rcvd[2*from[d]+dir]++;
    q[d].dst = (char *)rb[d]; q[d].src = MB.mine + mb_slot_off( from[d], dir, k&1 ); q[d].nbytes = (unsigned)rn[d]; q[d].seq = k+1;
    q[d].flag = reinterpret_cast<const unsigned *>( MB.mine + mb_flag_off( W, from[d], dir, k&1 ) );
  }
  const size_t big = std::max( std::max( nslo, nshi ), std::max( nrlo, nrhi ) );
  const int nb = (int)std::min<size_t>( 64, std::max<size_t>( 1, big/( 256*16*4 ) ) );
  if( p[0].flag || p[1].flag ) { k_put<<< dim3( nb, 2 ), 256, 0, r.stream >>>( p[0], p[1], MB.d_done ); VPB_CHECK_LAUNCH(); r.launches++; }
  if( q[0].flag || q[1].flag ) { k_get<<< dim3( nb, 2 ), 256, 0, r.stream >>>( q[0], q[1], MB.d_err, put_timeout_clocks() ); VPB_CHECK_LAUNCH(); r.launches++; }
  return true;
}

} // namespace

// ---- fused exchanges: the caller's own kernels write into / read from the mailboxes ----
// comm_link_reserve hands out, for one exchange with (rank_lo, rank_hi), the
// mailbox slots that put_sendrecv would have used: where this rank's "lo" and
// "hi" messages go in the neighbours' memory (dst, flag to publish, sequence
// number) and where the neighbours' messages arrive in ours (src, flag to
// wait for, sequence number).  The pack kernel stores the plane straight into
// dst over NVLink and publishes; the unpack kernel acquires the flag and reads
// src.  Returns false (nothing reserved) when the put transport is off, the
// message does not fit a slot, or VPIC_B200_TRANSPORT=put-unfused.
bool comm_link_reserve( int rank_lo, int rank_hi, size_t nbytes, CommLink * out ) {
  static int unfused = -1;
  if( unfused<0 ) { const char * e = getenv( "VPIC_B200_TRANSPORT" ); unfused = ( e && !strcmp( e, "put-unfused" ) ) ? 1 : 0; }
  *out = CommLink();
  if( !MB.on && !MB.tried && g_world_size>1 ) need_comm( "a face exchange" );   // an unmodified deck's first exchange bootstraps the transport
  if( !MB.on || unfused || nbytes>MB_SLOT ) return false;
  const int W = g_world_size, me = g_world_rank;
  const int peer[2] = { rank_lo, rank_hi };
  for( int d=0; d<2; d++ ) if( peer[d]>=0 ) {
    const unsigned k = MB.sent[2*peer[d]+d]++;
    char * box = mailbox_of( peer[d] );
    out->dst[d] = box + mb_slot_off( me, d, k&1 );
    out->pub_flag[d] = reinterpret_cast<unsigned *>( box + mb_flag_off( W, me, d, k&1 ) );
    out->pub_seq[d] = k+1;
    const int dir = 1-d;                       // what arrives from rank_lo was its "hi" message
    const unsigned j = MB.rcvd[2*peer[d]+dir]++;
    out->src[d] = MB.mine + mb_slot_off( peer[d], dir, j&1 );
    out->wait_flag[d] = reinterpret_cast<const unsigned *>( MB.mine + mb_flag_off( W, peer[d], dir, j&1 ) );
    out->wait_seq[d] = j+1;
  }
  if( !MB.d_done_many ) { VPB_CUDA( cudaMalloc( &MB.d_done_many, sizeof(unsigned)*64 ) ); VPB_CUDA( cudaMemset( MB.d_done_many, 0, sizeof(unsigned)*64 ) ); }
  out->done = MB.d_done_many;
  return true;
}

// Pairwise exchange along one axis.  Messages to the same peer are matched in
// issue order, so when both neighbours are the same rank (two ranks, periodic)
// the peer's "low" message must be the first thing we receive from it, and it
// lands on our HIGH side: sends go lo,hi and receives go hi,lo on every rank.
void comm_sendrecv( int rank_lo, int rank_hi,
                    const void * slo, size_t nslo, const void * shi, size_t nshi,
                    void * rlo, size_t nrlo, void * rhi, size_t nrhi ) {
  need_comm( "a face exchange" );
  Runtime & r = rt();
  // Transport per NEIGHBOUR PAIR, from the two sizes both ends of the pair know
  // (what we send them, what they send us): a message that does not fit a
  // mailbox slot takes that pair to NCCL without the other neighbour having to
  // know about it.
  if( MB.on ) {
    const bool put_lo = rank_lo>=0 && nslo<=MB_SLOT && nrlo<=MB_SLOT, put_hi = rank_hi>=0 && nshi<=MB_SLOT && nrhi<=MB_SLOT;
    put_sendrecv( put_lo ? rank_lo : -1, put_hi ? rank_hi : -1, slo, nslo, shi, nshi, rlo, nrlo, rhi, nrhi );
    if( put_lo ) rank_lo = -1;
    if( put_hi ) rank_hi = -1;
    if( rank_lo<0 && rank_hi<0 ) return;
  }
  if( !N.comm ) VPB_ERROR( "a %zu-byte message does not fit a mailbox slot and there is no NCCL communicator to carry it",
                           std::max( std::max( nslo, nshi ), std::max( nrlo, nrhi ) ) );
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
  if( n>RED_MAX ) VPB_ERROR( "allreduce of %d doubles", n );
  if( MB.on ) {
    const int W = g_world_size;
    if( !MB.d_peers ) {
      std::vector<char *> h( W );
      for( int k=0; k<W; k++ ) h[k] = mailbox_of( k );
      VPB_CUDA( cudaMalloc( &MB.d_peers, sizeof(char *)*W ) );
      VPB_CUDA( cudaMemcpy( MB.d_peers, h.data(), sizeof(char *)*W, cudaMemcpyHostToDevice ) );
    }
    const unsigned k = MB.red_seq++;
    h2d( MB.d_red, v, sizeof(double)*(size_t)n, r.stream );
    k_allsum_put<<< 1, 256, 0, r.stream >>>( MB.d_peers, W, g_world_rank, (int)( k&1 ), k+1, MB.d_red, n );
    k_allsum_get<<< 1, RED_MAX, 0, r.stream >>>( MB.mine, W, (int)( k&1 ), k+1, MB.d_red, n, MB.d_err, put_timeout_clocks() );
    VPB_CHECK_LAUNCH(); r.launches += 2;
    d2h( v, MB.d_red, sizeof(double)*(size_t)n, r.stream );
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    return;
  }
  if( !N.d_red ) VPB_CUDA( cudaMalloc( &N.d_red, sizeof(double)*64 ) );
  h2d( N.d_red, v, sizeof(double)*(size_t)n, r.stream );
  VPB_NCCL( N.AllReduce( N.d_red, N.d_red, (size_t)n, ncclFloat64, ncclSum, N.comm, r.stream ) );
  d2h( v, N.d_red, sizeof(double)*(size_t)n, r.stream );
  VPB_CUDA( cudaStreamSynchronize( r.stream ) );
  r.launches++;
}

bool comm_ready() { return N.comm!=nullptr || MB.on; }

} // namespace vpb

// src/util/mp/mp.h:76-79 of a host program built from the reference (weak)
extern "C" void mp_allgather_i( int * sbuf, int * rbuf, int n ) __attribute__((weak));

namespace vpb {

// Unmodified decks never call vpic_b200_comm_init.  When the host program has
// the reference's message-passing layer, the first exchange that needs a
// neighbour (collective by construction: every rank reaches its first ghost
// exchange) creates the NCCL id on rank 0 and ships it with the host
// program's own mp_allgather_i.
void bootstrap_from_host_program() {
  if( N.comm || MB.on || g_world_size<2 || !mp_allgather_i ) return;
  ensure_init();
  const int W = g_world_size;
  const char * tr = getenv( "VPIC_B200_TRANSPORT" );
  bool want_nccl = tr && !strcmp( tr, "nccl" );
  if( !want_nccl ) {
    // the mailboxes' IPC handles travel over the host program's allgather: no NCCL needed at all
    constexpr int NI = (int)( MB_REC/sizeof(int) );
    static_assert( MB_REC%sizeof(int)==0, "record is a whole number of ints" );
    std::vector<int> mine( NI, 0 ), all( (size_t)NI*W, 0 );
    const int ok = mailbox_alloc() ? 1 : 0;
    mine[0] = ok; memcpy( (char *)mine.data()+16, &MB.my_handle, sizeof(MB.my_handle) );
    mp_allgather_i( mine.data(), all.data(), NI );
    MB.tried = true;
    mailbox_install( (const char *)all.data() );
    if( MB.on ) return;
    want_nccl = true;
  }
  load_nccl();
  constexpr int NI = (int)( ( sizeof(ncclUniqueId) + sizeof(int) - 1 )/sizeof(int) );
  std::vector<int> mine( NI, 0 ), all( (size_t)NI*W, 0 );
  if( g_world_rank==0 ) { ncclUniqueId id; VPB_NCCL( N.GetUniqueId( &id ) ); memcpy( mine.data(), &id, sizeof(id) ); }
  mp_allgather_i( mine.data(), all.data(), NI );
  ncclUniqueId id; memcpy( &id, all.data(), sizeof(id) );
  VPB_NCCL( N.CommInitRank( &N.comm, W, id, g_world_rank ) );
  mailbox_setup();
}

void comm_shutdown() {
  if( N.comm ) { N.CommDestroy( N.comm ); N.comm = nullptr; }
  if( N.d_red ) { cudaFree( N.d_red ); N.d_red = nullptr; }
  for( char * q : MB.peer ) if( q ) cudaIpcCloseMemHandle( q );
  if( MB.mine ) cudaFree( MB.mine );
  if( MB.d_done ) cudaFree( MB.d_done );
  if( MB.d_done_many ) cudaFree( MB.d_done_many );
  if( MB.d_err ) cudaFree( MB.d_err );
  if( MB.d_red ) cudaFree( MB.d_red );
  if( MB.d_peers ) cudaFree( MB.d_peers );
  MB = Mailbox();
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
  mailbox_setup();
}

// Bootstrap WITHOUT NCCL: every rank asks for its mailbox record (80 bytes: export status + cudaIpc
// handle), the host program all-gathers the records by whatever means it has, every rank installs
// the full set.  Face exchanges, particle migration and the all-reduce then run entirely on our put
// kernels over peer memory -- also when several ranks share one GPU, which NCCL refuses.
int vpic_b200_comm_mailbox_record( void * out, int nbytes ) {
  if( !out || nbytes<(int)MB_REC ) VPB_ERROR( "Bad args" );
  ensure_init();
  const int ok = mailbox_alloc() ? 1 : 0;
  memset( out, 0, MB_REC ); memcpy( out, &ok, sizeof(int) ); memcpy( (char *)out+16, &MB.my_handle, sizeof(MB.my_handle) );
  return (int)MB_REC;
}

void vpic_b200_comm_init_handles( const void * all_records, int nbytes_each ) {
  if( !all_records || nbytes_each!=(int)MB_REC ) VPB_ERROR( "Bad args" );
  ensure_init();
  if( MB.on ) return;
  mailbox_alloc();
  MB.tried = true;
  mailbox_install( (const char *)all_records );
  if( !MB.on ) VPB_ERROR( "the mailbox of some rank could not be exported with cudaIpc: use vpic_b200_comm_init (NCCL)" );
}

// 1 = face exchanges and particle migration go through the put transport
// (NVLink peer-memory mailboxes), 0 = through ncclSend/ncclRecv.
int vpic_b200_comm_transport( void ) { return MB.on ? 1 : 0; }

int vpic_b200_comm_ready( void ) { return comm_ready() ? 1 : 0; }

} // extern "C"
