// vpb_sf.cu -- species<->field interface kernels on the device:
//   load_interpolator_array  (reference: src/sf_interface/pipeline/interpolator_array_pipeline.cc:20-126)
//   clear_accumulator_array  (clear_accumulators_pipeline.c:9-61)
//   reduce_accumulator_array (reduce_accumulators_pipeline.cc:14-243)  -> no-op, one device copy
//   unload_accumulator_array (unload_accumulator_pipeline.cc:17-137)
// All are streaming HBM-bound kernels; this file is compiled with -fmad=false
// so every expression rounds exactly like the reference's scalar build.
#include "vpb_runtime.h"

namespace vpb {

// Algorithmic traffic: 24 B read (6 planes) + 72 B written per voxel (96 B with the device padding).
// One thread per interior voxel; x is the fastest index so the six plane reads
// are unit-stride and the +y/+z neighbours come out of L1/L2.
__global__ void __launch_bounds__(256)
k_load_interpolator( float * __restrict__ ip, const float * __restrict__ f, GridDev g ) {
  // flat grid, blockIdx.y = z (see k_advance_b): blocks walk the planes in order.
  // A thread's 96 B record goes to shared memory first (rows of 112 B: the eight
  // lanes of a quarter warp hit eight different bank groups) and the warp then
  // writes its 32 records -- 3 KB, contiguous along x -- with full 512 B wavefronts
  // instead of six 16 B pieces per thread at a 96 B stride.
  constexpr int ROW = 7;                                   // float4 per staged record
  __shared__ float4 s_rec[8][32*ROW];
  __shared__ int    s_vox[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
  const bool live = ixy<g.nx*g.ny;
  int v = -1;
  if( live ) {
    const int z = blockIdx.y, y = ixy / g.nx, x = ixy - y*g.nx;
    v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
    const size_t P = (size_t)g.nvp;
    const float * ex = f + F_EX*P, * ey = f + F_EY*P, * ez = f + F_EZ*P;
    const float * bx = f + F_CBX*P, * by = f + F_CBY*P, * bz = f + F_CBZ*P;
    const float fourth = 0.25f, half = 0.5f;
    float w0, w1, w2, w3;
    float4 o0, o1, o2, o3, o4;
    w0 = ex[v]; w1 = ex[v+g.sy]; w2 = ex[v+g.sz]; w3 = ex[v+g.sy+g.sz];
    o0.x = fourth*( (w3 + w0) + (w1 + w2) );
    o0.y = fourth*( (w3 - w0) + (w1 - w2) );
    o0.z = fourth*( (w3 - w0) - (w1 - w2) );
    o0.w = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = ey[v]; w1 = ey[v+g.sz]; w2 = ey[v+1]; w3 = ey[v+g.sz+1];
    o1.x = fourth*( (w3 + w0) + (w1 + w2) );
    o1.y = fourth*( (w3 - w0) + (w1 - w2) );
    o1.z = fourth*( (w3 - w0) - (w1 - w2) );
    o1.w = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = ez[v]; w1 = ez[v+1]; w2 = ez[v+g.sy]; w3 = ez[v+1+g.sy];
    o2.x = fourth*( (w3 + w0) + (w1 + w2) );
    o2.y = fourth*( (w3 - w0) + (w1 - w2) );
    o2.z = fourth*( (w3 - w0) - (w1 - w2) );
    o2.w = fourth*( (w3 + w0) - (w1 + w2) );
    w0 = bx[v]; w1 = bx[v+1];    o3.x = half*( w1 + w0 ); o3.y = half*( w1 - w0 );
    w0 = by[v]; w1 = by[v+g.sy]; o3.z = half*( w1 + w0 ); o3.w = half*( w1 - w0 );
    w0 = bz[v]; w1 = bz[v+g.sz]; o4.x = half*( w1 + w0 ); o4.y = half*( w1 - w0 );
    o4.z = 0.f; o4.w = 0.f;
    float4 * row = s_rec[w] + lane*ROW;
    row[0] = o0; row[1] = o1; row[2] = o2; row[3] = o3; row[4] = o4; row[5] = make_float4( 0.f, 0.f, 0.f, 0.f );
  }
  s_vox[w][lane] = v;
  __syncwarp();
# pragma unroll
  for( int k=0; k<6; k++ ) {
    const int e = k*32 + lane, rec = e/6, part = e - rec*6;
    const int rv = s_vox[w][rec];
    if( rv>=0 ) reinterpret_cast<float4 *>( ip + (size_t)rv*IP_STRIDE )[part] = s_rec[w][rec*ROW + part];
  }
}

// jf += c*(four quarter-face sums); optionally fused with clear_jf (then jf = ...).
// Range x:1..nx+1, y:1..ny+1, z:1..nz+1.  48 B gathered + 12 B written per voxel.
template<bool CLEAR_FIRST>
__global__ void __launch_bounds__(256)
k_unload_accumulator( float * __restrict__ f, const float * __restrict__ a, GridDev g, float cx, float cy, float cz ) {
  const int mx = g.nx+1, my = g.ny+1;
  const size_t P = (size_t)g.nvp;
  float * jfx = f + F_JFX*P, * jfy = f + F_JFY*P, * jfz = f + F_JFZ*P;
  {
    const int ixy = blockIdx.x*blockDim.x + threadIdx.x;
    if( ixy>=mx*my ) return;
    const int z = blockIdx.y, y = ixy / mx, x = ixy - y*mx;
    const int v = (x+1) + g.sy*(y+1) + g.sz*(z+1);
    const float * a0  = a + (size_t)v*ACC_STRIDE;
    const float * ax  = a + (size_t)(v-1)*ACC_STRIDE;
    const float * ay  = a + (size_t)(v-g.sy)*ACC_STRIDE;
    const float * az  = a + (size_t)(v-g.sz)*ACC_STRIDE;
    const float * ayz = a + (size_t)(v-g.sy-g.sz)*ACC_STRIDE;
    const float * azx = a + (size_t)(v-g.sz-1)*ACC_STRIDE;
    const float * axy = a + (size_t)(v-1-g.sy)*ACC_STRIDE;
    const float sx_ = cx*( ( ( a0[0] + ay[1] ) + az[2] ) + ayz[3] );
    const float sy_ = cy*( ( ( a0[4] + az[5] ) + ax[6] ) + azx[7] );
    const float sz_ = cz*( ( ( a0[8] + ax[9] ) + ay[10] ) + axy[11] );
    if( CLEAR_FIRST ) { jfx[v] = 0.f + sx_; jfy[v] = 0.f + sy_; jfz[v] = 0.f + sz_; }
    else              { jfx[v] += sx_;     jfy[v] += sy_;      jfz[v] += sz_; }
  }
}


} // namespace vpb

using namespace vpb;

extern "C" {

void load_interpolator_array( vpb_interpolator_array_t * ia, const vpb_field_array_t * fa ) {
  if( !ia || !fa || ia->g!=fa->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  InterpMirror * mi = mirror( ia );
  FieldMirror  * mf = mirror( const_cast<vpb_field_array_t *>( fa ) );
  to_device( mf );
  const GridDev & g = mi->grid->dev;
  k_load_interpolator<<< dim3( div_up( (long long)g.nx*g.ny, 256 ), g.nz ), 256, 0, r.stream >>>( mi->d, mf->d, g );
  VPB_CHECK_LAUNCH(); r.launches++;
  finish_write( mi );
}

void clear_accumulator_array( vpb_accumulator_array_t * aa ) {
  if( !aa ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  AccMirror * m = mirror( aa );
  // One device copy; ghosts are never deposited into, so clearing the whole
  // array is equivalent to the reference's [VOXEL(1,1,1), VOXEL(nx,ny,nz)] range.
  VPB_CUDA( cudaMemsetAsync( m->d, 0, sizeof(float)*ACC_STRIDE*(size_t)m->grid->dev.nv, r.stream ) );
  if( m->d_priv ) VPB_CUDA( cudaMemsetAsync( m->d_priv, 0, sizeof(float)*ACC_STRIDE*(size_t)m->grid->dev.nv*(size_t)m->n_priv, r.stream ) );
  r.launches++;
  mark_dev_written( m );
  if( r.coherence==VPB_COHERENCE_STRICT && aa->a ) {
    // keep the host's copies (shared + one per pipeline) coherent with "all
    // clear" as well: the zeros come down from the device copy by DMA, which
    // is several times faster than one host thread's memset
    const int nx = aa->g->nx, ny = aa->g->ny, nz = aa->g->nz;
    int i0 = ( VPB_VOXEL(1,1,1,nx,ny,nz)/2 )*2;
    int n  = ( ( ( VPB_VOXEL(nx,ny,nz,nx,ny,nz) - i0 + 1 ) + 1 )/2 )*2;
    pin_host( aa->a, sizeof(vpb_accumulator_t)*( (size_t)aa->n_pipeline*aa->stride + (size_t)m->grid->dev.nv ) );
    for( int c=0; c<=aa->n_pipeline; c++ )
      d2h( aa->a + (size_t)c*aa->stride + i0, m->d + (size_t)ACC_STRIDE*i0, sizeof(vpb_accumulator_t)*(size_t)n, r.stream );
    VPB_CUDA( cudaStreamSynchronize( r.stream ) );
    m->host_valid = true;
  }
}

void reduce_accumulator_array( vpb_accumulator_array_t * aa ) {
  if( !aa ) VPB_ERROR( "Bad args" );
  // FAST: all device deposits already land in the single device copy (= host
  // copy 0).  PARITY: add the per-pipeline copies with the reference's tree.
  AccMirror * m = mirror( aa );
  if( rt().deposit_mode==VPB_DEPOSIT_PARITY && m->dev_valid ) {
    reduce_accumulators_parity( m );
    finish_write( m );
  }
}

void unload_accumulator_array( vpb_field_array_t * fa, const vpb_accumulator_array_t * aa ) {
  if( !fa || !aa || fa->g!=aa->g ) VPB_ERROR( "Bad args" );
  Runtime & r = rt();
  FieldMirror * mf = mirror( fa );
  AccMirror   * ma = mirror( const_cast<vpb_accumulator_array_t *>( aa ) );
  to_device( mf ); to_device( ma );
  const vpb_grid_t * hg = fa->g;
  // coefficients formed in double and rounded once (unload_accumulator_pipeline.cc:130-132)
  const float cx = 0.25 * hg->rdy * hg->rdz / hg->dt;
  const float cy = 0.25 * hg->rdz * hg->rdx / hg->dt;
  const float cz = 0.25 * hg->rdx * hg->rdy / hg->dt;
  const GridDev & g = mf->grid->dev;
  k_unload_accumulator<false><<< dim3( div_up( (long long)(g.nx+1)*(g.ny+1), 256 ), g.nz+1 ), 256, 0, r.stream >>>( mf->d, ma->d, g, cx, cy, cz );
  VPB_CHECK_LAUNCH(); r.launches++;
  finish_write( mf );
}

} // extern "C"
