/* oracle shim: <linux/bio.h> -> lz4e_kshim.h (test infrastructure) */
#include "lz4e_kshim.h"
