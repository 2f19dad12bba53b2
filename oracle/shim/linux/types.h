/* oracle shim: <linux/types.h> -> lz4e_kshim.h (test infrastructure) */
#include "lz4e_kshim.h"
