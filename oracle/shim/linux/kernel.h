/* oracle shim: <linux/kernel.h> -> lz4e_kshim.h (test infrastructure) */
#include "lz4e_kshim.h"
