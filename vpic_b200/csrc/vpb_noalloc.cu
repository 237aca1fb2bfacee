// vpb_noalloc.cu -- vpic_b200_energy_f_noalloc: the vacuum field-energy
// reduction (energy_f, vpb_fields.cu: k_energy_f<true>) for callers that may
// not allocate -- code under CUDA-graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The regular entry point (vpic_b200_energy_f) puts its per-block partial sums
// in the runtime's grow-on-demand scratch arena (cudaMalloc).  This variant
// keeps them in a FILE-SCOPE device array of fixed size, so it allocates
// nothing, ever: no cudaMalloc / cudaFree, no mirror look-up, no runtime
// state.  Inputs and outputs are the caller's device buffers.
//
//  * Limit: NA_MAX_BLOCKS partial rows of NA_CELLS_PER_BLOCK cells =
//    33 554 432 interior cells (exactly BASELINE configs[3]'s 512x256x256 box);
//    larger grids are rejected with an error code and a message.  The array
//    is 384 KB and every GPU that loads the library holds it until exit.
//  * The array is shared by every call and keeps the last call's contents.
//    Nothing in it has to start at zero: a call overwrites every row it later
//    reads (rows [0, nblock)) and reads no other.  Calls that OVERLAP on the
//    GPU (two streams, or two host threads) race on it: serialise them.
//  * In host code `g_na_partial` is a host-side stand-in, not a device
//    address; the address the kernels get comes from cudaGetSymbolAddress,
//    looked up once per device with that device current.  The first call on a
//    device therefore has to happen outside stream capture.
#include <cuda_runtime.h>
#include <cstdio>
#include "vpic_b200/vpic_b200.h"

namespace {

constexpr int NA_BLOCK           = 256;                  // threads per block
constexpr int NA_CELLS_PER_BLOCK = 4096;                 // 16 cells per thread
constexpr int NA_MAX_BLOCKS      = 8192;                 // rows of the static scratch
constexpr long long NA_MAX_CELLS = (long long)NA_MAX_BLOCKS*NA_CELLS_PER_BLOCK;   // 33 554 432
constexpr int NA_MAX_DEVICES     = 64;

static __device__ __align__(256) double g_na_partial[ 6*NA_MAX_BLOCKS ];   // 393 216 B

// Same per-cell expressions and float -> double promotion as k_energy_f<true>
// (vacuum_energy_f_pipeline.h:18-73); only the assignment of cells to blocks
// differs (fixed 4096-cell rows instead of a grid-stride loop), so the double
// sums associate differently: results agree to ~1e-15 relative.
__global__ void __launch_bounds__( NA_BLOCK )
k_energy_f_noalloc( double * __restrict__ partial, const float * __restrict__ f,
                    int nx, int ny, int nz, int sy, int sz, size_t P,
                    float qepsx, float qepsy, float qepsz, float hrmux, float hrmuy, float hrmuz ) {
  const int ncell = nx*ny*nz;
  const float * ex = f + 0*P, * ey = f + 1*P, * ez = f + 2*P;      // F_EX, F_EY, F_EZ
  const float * bx = f + 4*P, * by = f + 5*P, * bz = f + 6*P;      // F_CBX, F_CBY, F_CBZ
  double e[6] = {0,0,0,0,0,0};
  const int first = blockIdx.x*NA_CELLS_PER_BLOCK;
  for( int k=threadIdx.x; k<NA_CELLS_PER_BLOCK; k+=NA_BLOCK ) {
    const int i = first + k;
    if( i>=ncell ) break;
    const int x = i % nx, t = i / nx, y = t % ny, z = t / ny;
    const int v = (x+1) + sy*(y+1) + sz*(z+1);
    const int vx = v+1, vy = v+sy, vz = v+sz, vyz = v+sy+sz, vzx = v+sz+1, vxy = v+1+sy;
    e[0] += (double)( qepsx*( ex[v]*ex[v] + ex[vy]*ex[vy] + ex[vz]*ex[vz] + ex[vyz]*ex[vyz] ) );
    e[1] += (double)( qepsy*( ey[v]*ey[v] + ey[vz]*ey[vz] + ey[vx]*ey[vx] + ey[vzx]*ey[vzx] ) );
    e[2] += (double)( qepsz*( ez[v]*ez[v] + ez[vx]*ez[vx] + ez[vy]*ez[vy] + ez[vxy]*ez[vxy] ) );
    e[3] += (double)( hrmux*( bx[v]*bx[v] + bx[vx]*bx[vx] ) );
    e[4] += (double)( hrmuy*( by[v]*by[v] + by[vy]*by[vy] ) );
    e[5] += (double)( hrmuz*( bz[v]*bz[v] + bz[vz]*bz[vz] ) );
  }
  __shared__ double s[6][NA_BLOCK/32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
# pragma unroll
  for( int c=0; c<6; c++ ) {
    double vsum = e[c];
    for( int o=16; o; o>>=1 ) vsum += __shfl_down_sync( 0xffffffffu, vsum, o );
    if( lane==0 ) s[c][w] = vsum;
  }
  __syncthreads();
  if( threadIdx.x<6 ) {
    double vsum = 0;
    for( int k=0; k<NA_BLOCK/32; k++ ) vsum += s[threadIdx.x][k];
    partial[ (size_t)blockIdx.x*6 + threadIdx.x ] = vsum;       // every row in [0, gridDim.x) is overwritten
  }
}

// One block, warp c adds component c of rows [0, nblock) in a fixed order.
__global__ void __launch_bounds__( 6*32 )
k_energy_f_noalloc_sum( double * __restrict__ out, const double * __restrict__ partial, int nblock, double scale ) {
  const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;
  double vsum = 0;
  for( int b=lane; b<nblock; b+=32 ) vsum += partial[ (size_t)b*6 + c ];
  for( int o=16; o; o>>=1 ) vsum += __shfl_down_sync( 0xffffffffu, vsum, o );
  if( lane==0 ) out[c] = vsum*scale;
}

// Device address of g_na_partial on the current device (nullptr + message on failure).
double * na_scratch_ptr() {
  static double * ptr[ NA_MAX_DEVICES ] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice( &dev );
  if( e!=cudaSuccess || dev<0 || dev>=NA_MAX_DEVICES ) {
    fprintf( stderr, "vpic_b200_energy_f_noalloc: cudaGetDevice failed or device %d out of range: %s\n", dev, cudaGetErrorString( e ) );
    return nullptr;
  }
  if( !ptr[dev] ) {
    double * p = nullptr;
    e = cudaGetSymbolAddress( (void **)&p, g_na_partial );     // the symbol is a host-side handle; this is its device address
    if( e!=cudaSuccess || !p ) {
      fprintf( stderr, "vpic_b200_energy_f_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString( e ) );
      return nullptr;
    }
    ptr[dev] = p;
  }
  return ptr[dev];
}

} // namespace

extern "C" int vpic_b200_energy_f_noalloc( double * d_en6, const float * d_field_soa, int nx, int ny, int nz,
                                           const float * eps_rmu6, double scale, void * stream ) {
  if( !d_en6 || !d_field_soa || !eps_rmu6 || nx<1 || ny<1 || nz<1 ) {
    fprintf( stderr, "vpic_b200_energy_f_noalloc: Bad args\n" );
    return -1;
  }
  const long long ncell = (long long)nx*ny*nz;
  if( ncell>NA_MAX_CELLS ) {
    fprintf( stderr, "vpic_b200_energy_f_noalloc: %d x %d x %d = %lld cells exceed the static scratch (at most %lld cells); "
                     "use vpic_b200_energy_f\n", nx, ny, nz, ncell, NA_MAX_CELLS );
    return -2;
  }
  double * partial = na_scratch_ptr();
  if( !partial ) return -3;
  const int sy = nx+2, sz = ( nx+2 )*( ny+2 );
  const long long nv = (long long)sz*( nz+2 );
  const size_t P = (size_t)( ( nv + 31 ) & ~31ll );                      // plane pitch of the SoA field mirror (vpb_runtime.cu: nvp)
  const int nblock = (int)( ( ncell + NA_CELLS_PER_BLOCK - 1 )/NA_CELLS_PER_BLOCK );
  cudaStream_t st = (cudaStream_t)stream;
  k_energy_f_noalloc<<< nblock, NA_BLOCK, 0, st >>>( partial, d_field_soa, nx, ny, nz, sy, sz, P,
                                                     0.25f*eps_rmu6[0], 0.25f*eps_rmu6[1], 0.25f*eps_rmu6[2],
                                                     0.50f*eps_rmu6[3], 0.50f*eps_rmu6[4], 0.50f*eps_rmu6[5] );
  k_energy_f_noalloc_sum<<< 1, 6*32, 0, st >>>( d_en6, partial, nblock, scale );
  const cudaError_t e = cudaGetLastError();
  if( e!=cudaSuccess ) {
    fprintf( stderr, "vpic_b200_energy_f_noalloc: launch failed: %s\n", cudaGetErrorString( e ) );
    return -3;
  }
  return 0;
}
