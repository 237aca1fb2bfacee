/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of boundary_p (F14) and accumulate_rhob (F15). */
#include "orc_common.h"

/* F15: rho_p.cc:125-171 -- filled in below. */
void orc_accumulate_rhob( vpb_field_t * f, const vpb_particle_t * p, const vpb_grid_t * g, float qsp ) {
  (void)f; (void)p; (void)g; (void)qsp;
  ORC_FAIL( "accumulate_rhob: not restated yet" );
}

/* F14: boundary_p.cc:36-518.  Round-1 state: with every face periodic-to-self
   or reflecting, move_p never leaves a mover behind (boundary_p.cc:111-112),
   so the only legal state is "no movers". */
void orc_boundary_p( vpb_species_t * sp_list, vpb_field_array_t * fa, vpb_accumulator_array_t * aa ) {
  if( !sp_list || !fa || !aa ) ORC_FAIL( "Bad args" );
  for( vpb_species_t * sp=sp_list; sp; sp=sp->next )
    if( sp->nm>0 ) ORC_FAIL( "boundary_p with pending movers: not restated yet" );
}
