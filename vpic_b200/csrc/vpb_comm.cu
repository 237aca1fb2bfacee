// vpb_comm.cu -- multi-GPU transport for face exchanges and reductions.
// Reference counterpart: src/util/mp (MPI ports) as used by
// src/field_advance/standard/remote.c and src/boundary/boundary_p.cc.
// Round-1 state: single-GPU builds route periodic faces to themselves inside
// vpb_fields.cu; a rank with a REMOTE neighbour needs the NCCL transport, which
// is installed by vpic_b200_comm_init (not yet available) -- until then these
// entry points fail loudly instead of silently computing something else.
#include "vpb_runtime.h"

namespace vpb {

void comm_exchange_faces( FieldMirror *, int X, float *, float *, float *, float *, int ) {
  VPB_ERROR( "face exchange along axis %d needs a remote neighbour but no multi-GPU transport is initialised", X );
}

void comm_allsum_d( double *, int ) {
  VPB_ERROR( "allreduce requested with world_size=%d but no multi-GPU transport is initialised", g_world_size );
}

} // namespace vpb
