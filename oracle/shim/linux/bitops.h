/* oracle shim: <linux/bitops.h> -> lz4e_kshim.h (test infrastructure) */
#include "lz4e_kshim.h"
