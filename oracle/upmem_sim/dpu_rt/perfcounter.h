/* DPU runtime header stand-in: see upmem_sim_dpu_rt.h */
#include "upmem_sim_dpu_rt.h"
