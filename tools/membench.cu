// Micro-benchmark: how fast can a B200 stream S read planes + W write planes
// (one float or one float4 per thread and plane)?  Sizes like the 256^3 Yee planes.
#include <cstdio>
#include <cuda_runtime.h>
template<int R, int W, typename T>
__global__ void __launch_bounds__(256) k_stream( const T * __restrict__ in, T * __restrict__ out, size_t n, size_t pitch ) {
  const size_t stride = (size_t)gridDim.x*blockDim.x;
  for( size_t i=(size_t)blockIdx.x*blockDim.x+threadIdx.x; i<n; i+=stride ) {
    T v[R];
#   pragma unroll
    for( int r=0; r<R; r++ ) v[r] = in[ r*pitch + i ];
    float acc = 0;
#   pragma unroll
    for( int r=0; r<R; r++ ) acc += *(float *)&v[r];
#   pragma unroll
    for( int w=0; w<W; w++ ) { T o = v[w%R]; *(float *)&o += acc; out[ w*pitch + i ] = o; }
  }
}
template<int R, int W, typename T> void run( const char * name, void * in, void * out, size_t nfloat, size_t pitchf, int blocks ) {
  const size_t n = nfloat/( sizeof(T)/4 ), pitch = pitchf/( sizeof(T)/4 );
  cudaEvent_t a, b; cudaEventCreate( &a ); cudaEventCreate( &b );
  for( int k=0; k<3; k++ ) k_stream<R,W,T><<< blocks, 256 >>>( (const T *)in, (T *)out, n, pitch );
  cudaEventRecord( a );
  for( int k=0; k<20; k++ ) k_stream<R,W,T><<< blocks, 256 >>>( (const T *)in, (T *)out, n, pitch );
  cudaEventRecord( b ); cudaEventSynchronize( b );
  float ms; cudaEventElapsedTime( &ms, a, b ); ms /= 20;
  printf( "%-28s blocks %6d  %.3f ms  %.0f GB/s\n", name, blocks, ms, (double)( R+W )*nfloat*4/ms/1e6 );
}
int main() {
  const size_t nf = 17173512, pitch = ( nf + 31 ) & ~(size_t)31;
  void * in, * out;
  cudaMalloc( &in, pitch*4*12 ); cudaMalloc( &out, pitch*4*12 );
  cudaMemset( in, 0, pitch*4*12 ); cudaMemset( out, 0, pitch*4*12 );
  for( int blocks : { 148*8, 148*16, 148*32, 67084 } ) {
    run<1,1,float >( "1R+1W float",  in, out, nf, pitch, blocks );
    run<1,1,float4>( "1R+1W float4", in, out, nf & ~(size_t)3, pitch, blocks );
    run<6,3,float >( "6R+3W float",  in, out, nf, pitch, blocks );
    run<6,3,float4>( "6R+3W float4", in, out, nf & ~(size_t)3, pitch, blocks );
    run<9,6,float >( "9R+6W float",  in, out, nf, pitch, blocks );
    run<9,6,float4>( "9R+6W float4", in, out, nf & ~(size_t)3, pitch, blocks );
  }
  return 0;
}
