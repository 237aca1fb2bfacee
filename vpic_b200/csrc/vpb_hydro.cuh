// vpb_hydro.cuh -- per-particle part of accumulate_hydro_p
// (src/species_advance/standard/hydro_p.c:62-137), shared by the atomic kernel
// (vpb_hydro.cu) and the reference-order kernel (vpb_parity.cu).
#pragma once
#include "vpb_push.cuh"

namespace vpb {

constexpr int HY_STRIDE = 16;   // floats per node, device layout == host hydro_t (64 B)

struct HydroCoef { float c, qsp, mspc, qdt_2mc, qdt_4mc2, r8V; };

struct HydroParticle {
  float wn[8];                  // (1/8)(w/V) trilinear weights, node order (0,0,0) (1,0,0) (0,1,0) (1,1,0) (0,0,1) ...
  float ux, uy, uz, vx, vy, vz, ke_mc;
};

// Half a Boris step (the momentum is half a step behind the position), the
// velocity, c(gamma-1) and the eight weights.  hydro_p.c:91 calls the double
// sqrt on a float and rounds back, which equals sqrtf; :96 is a double
// expression because of its (1./3.) and 0.4 constants.
__device__ __forceinline__ HydroParticle hydro_particle( float4 r0, float4 r1, const Interp & f, const HydroCoef & k ) {
  HydroParticle h;
  float dx = r0.x, dy = r0.y, dz = r0.z;
  float ux = r1.x, uy = r1.y, uz = r1.z;
  ux += k.qdt_2mc*( ( f.ex + dy*f.dexdy ) + dz*( f.dexdz + dy*f.d2exdydz ) );
  uy += k.qdt_2mc*( ( f.ey + dz*f.deydz ) + dx*( f.deydx + dz*f.d2eydzdx ) );
  uz += k.qdt_2mc*( ( f.ez + dx*f.dezdx ) + dy*( f.dezdy + dx*f.d2ezdxdy ) );
  float w5 = f.cbx + dx*f.dcbxdx, w6 = f.cby + dy*f.dcbydy, w7 = f.cbz + dz*f.dcbzdz;
  float ke_mc = ux*ux + uy*uy + uz*uz;
  float vz = sqrtf( 1.0f+ke_mc );
  ke_mc *= k.c/( vz+1.0f );
  vz = k.c/vz;
  float w0 = k.qdt_4mc2*vz;
  float w1 = w5*w5 + w6*w6 + w7*w7;
  float w2 = w0*w0*w1;
  float w3 = (float)( (double)w0*( 1.0 + ( 1./3. )*(double)w2*( 1.0 + 0.4*(double)w2 ) ) );
  float w4 = w3/( 1.0f + w1*w3*w3 ); w4 += w4;
  w0 = ux + w3*( uy*w7 - uz*w6 );
  w1 = uy + w3*( uz*w5 - ux*w7 );
  w2 = uz + w3*( ux*w6 - uy*w5 );
  ux += w4*( w1*w7 - w2*w6 );
  uy += w4*( w2*w5 - w0*w7 );
  uz += w4*( w0*w6 - w1*w5 );
  h.ux = ux; h.uy = uy; h.uz = uz;
  h.vx = ux*vz; h.vy = uy*vz; h.vz = vz*uz; h.ke_mc = ke_mc;
  w0 = k.r8V*r1.w; dx *= w0; w1 = w0+dx; w0 -= dx;
  w3 = 1.0f+dy; w2 = w0*w3; w3 *= w1; dy = 1.0f-dy; w0 *= dy; w1 *= dy;
  w7 = 1.0f+dz; w4 = w0*w7; w5 = w1*w7; w6 = w2*w7; w7 *= w3;
  dz = 1.0f-dz; w0 *= dz; w1 *= dz; w2 *= dz; w3 *= dz;
  h.wn[0] = w0; h.wn[1] = w1; h.wn[2] = w2; h.wn[3] = w3; h.wn[4] = w4; h.wn[5] = w5; h.wn[6] = w6; h.wn[7] = w7;
  return h;
}

// The 14 increments of one node (ACCUM_HYDRO, hydro_p.c:139-157), hydro_t order.
__device__ __forceinline__ void hydro_node_terms( float * __restrict__ o, float wn, const HydroCoef & k,
                                                  float ux, float uy, float uz, float vx, float vy, float vz, float ke_mc ) {
  float t = k.qsp*wn;
  o[0] = t*vx; o[1] = t*vy; o[2] = t*vz; o[3] = t;
  t = k.mspc*wn;
  const float px = t*ux, py = t*uy, pz = t*uz;
  o[4] = px; o[5] = py; o[6] = pz; o[7] = t*ke_mc;
  o[8] = px*vx; o[9] = py*vy; o[10] = pz*vz;
  o[11] = py*vz; o[12] = pz*vx; o[13] = px*vy;
}

} // namespace vpb
