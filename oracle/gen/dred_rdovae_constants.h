/* Stand-in: only the two maxima nnet.c reads (nnet.c:324, 450). */
#ifndef DRED_RDOVAE_CONSTANTS_H
#define DRED_RDOVAE_CONSTANTS_H
#define DRED_MAX_RNN_NEURONS 96
#define DRED_MAX_CONV_INPUTS 128
#endif
