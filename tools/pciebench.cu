// PCIe rates the STRICT-coherence path can count on: page-locked posix_memalign
// memory (what the reference's arrays become after cudaHostRegister), h2d, d2h
// and both at once on two streams, for several chunk sizes.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
int main() {
  const size_t N = (size_t)4<<30;
  void * h0, * h1; posix_memalign( &h0, 4096, N ); posix_memalign( &h1, 4096, N );
  for( size_t i=0; i<N; i+=4096 ) { ((char *)h0)[i] = 1; ((char *)h1)[i] = 1; }
  cudaHostRegister( h0, N, cudaHostRegisterDefault ); cudaHostRegister( h1, N, cudaHostRegisterDefault );
  void * d0, * d1; cudaMalloc( &d0, N ); cudaMalloc( &d1, N );
  cudaStream_t s0, s1; cudaStreamCreate( &s0 ); cudaStreamCreate( &s1 );
  cudaEvent_t a, b; cudaEventCreate( &a ); cudaEventCreate( &b );
  for( size_t chunk : { (size_t)8<<20, (size_t)64<<20, (size_t)512<<20, N } ) {
    for( int mode=0; mode<3; mode++ ) {
      cudaDeviceSynchronize();
      cudaEventRecord( a, s0 );
      cudaStreamWaitEvent( s1, a, 0 );
      for( size_t o=0; o<N; o+=chunk ) {
        if( mode!=1 ) cudaMemcpyAsync( (char *)d0+o, (char *)h0+o, chunk, cudaMemcpyHostToDevice, s0 );
        if( mode!=0 ) cudaMemcpyAsync( (char *)h1+o, (char *)d1+o, chunk, cudaMemcpyDeviceToHost, s1 );
      }
      cudaEventRecord( b, s1 ); cudaStreamWaitEvent( s0, b, 0 );
      cudaEventRecord( b, s0 ); cudaEventSynchronize( b );
      float ms; cudaEventElapsedTime( &ms, a, b );
      printf( "chunk %5zu MB  %-10s %.1f ms  %.1f GB/s per direction\n", chunk>>20, mode==0 ? "h2d" : mode==1 ? "d2h" : "both", ms, N/ms/1e6 );
    }
  }
  return 0;
}
