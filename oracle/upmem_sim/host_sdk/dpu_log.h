/* <dpu_log.h> stand-in: the reference only includes it (dpu_log_read call is commented out). */
#include "dpu.h"
