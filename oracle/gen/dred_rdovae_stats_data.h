/* Stand-in for the generated dred_rdovae_stats_data.h: the decoder network does not read the statistical model. */
#ifndef DRED_RDOVAE_STATS_DATA_H
#define DRED_RDOVAE_STATS_DATA_H
#include "opus_types.h"
#endif
