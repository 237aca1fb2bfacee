/* TEST INFRASTRUCTURE -- NOT PART OF THE PRODUCT.
 * Oracle restatement of boundary_p (F14) and accumulate_rhob (F15). */
#include "orc_common.h"

/* F15: rho_p.cc:125-171 -- filled in below. */
