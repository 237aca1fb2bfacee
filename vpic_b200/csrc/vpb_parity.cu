// vpb_parity.cu -- PARITY deposit mode of advance_p: currents are summed in the
// reference's order (particle array order, advance_p_pipeline.cc:54,80-82) so
// the accumulator is bit-identical to the scalar reference build.
#include "vpb_push.cuh"

namespace vpb {

void advance_p_parity( SpeciesMirror *, AccMirror *, InterpMirror *, const PushCoef & ) {
  VPB_ERROR( "VPB_DEPOSIT_PARITY is not available in this build yet" );
}

} // namespace vpb
