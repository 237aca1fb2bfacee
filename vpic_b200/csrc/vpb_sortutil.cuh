// vpb_sortutil.cuh -- warp-level helpers shared by sort_p and the PARITY deposit.
#pragma once
#include "vpb_runtime.h"

namespace vpb {

// FIX_SMEM ints of shared memory per warp for segments too long for registers;
// kept small so that 16 blocks fit on an SM (longer segments sort in global memory)
constexpr int FIX_SMEM = 512, FIX_WARPS = 4;

__device__ __forceinline__ void bitonic_mem( int * a, int n2, int lane ) {   // n2 power of two, padded with INT_MAX
  for( int k=2; k<=n2; k<<=1 ) for( int j=k>>1; j>0; j>>=1 ) {
    for( int t=lane; t<n2/2; t+=32 ) {
      const int lo = ( ( t & ~(j-1) )<<1 ) | ( t & (j-1) ), hi = lo | j;
      const bool up = ( lo & k )==0;
      const int x = a[lo], y = a[hi];
      if( ( x>y )==up ) { a[lo] = y; a[hi] = x; }
    }
    __syncwarp();
  }
}


} // namespace vpb
