// vpb_boundary.cu -- boundary_p: particles that left the rank's domain.
// Reference: src/boundary/boundary_p.cc:36-518.
#include "vpb_runtime.h"

using namespace vpb;

extern "C" void boundary_p( void * pbc_list, vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa ) {
  (void)pbc_list;
  if( !sp_list || !fa || !aa ) VPB_ERROR( "Bad args" );
  // On a single rank whose faces are periodic-to-self or reflecting, move_p
  // never hands a particle over (boundary_p.cc:111-112), so there is nothing
  // to do unless some species has movers.
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next ) {
    SpeciesMirror * m = mirror( sp );
    if( m->nm>0 ) VPB_ERROR( "species \"%s\" has %d particles on domain boundaries; device boundary_p "
                             "(absorb / remote migration) is not available in this build yet", sp->name, m->nm );
  }
}
