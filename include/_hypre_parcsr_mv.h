/* _hypre_parcsr_mv.h -- forwarding header: /root/reference/src/HypreSystem.h:18-24 includes hypre's headers by these names;
 * libb200hypre keeps its whole C ABI in HYPRE.h. */
#include "HYPRE.h"
