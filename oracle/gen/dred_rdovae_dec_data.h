/* Stand-in for the generated dred_rdovae_dec_data.h (torch/rdovae/export_rdovae_weights.py:178-222: the layers of
 * CoreDecoder, torch/rdovae/rdovae/rdovae.py:342-413, under the names src/dred_rdovae_dec.c uses). */
#ifndef DRED_RDOVAE_DEC_DATA_H
#define DRED_RDOVAE_DEC_DATA_H

#include "nnet.h"
#include "dred_rdovae_constants.h"

#define DEC_DENSE1_OUT_SIZE 256
#define DEC_DENSE2_OUT_SIZE 256
#define DEC_DENSE2_STATE_SIZE 256
#define DEC_DENSE3_OUT_SIZE 256
#define DEC_DENSE4_OUT_SIZE 256
#define DEC_DENSE4_STATE_SIZE 256
#define DEC_DENSE5_OUT_SIZE 256
#define DEC_DENSE6_OUT_SIZE 256
#define DEC_DENSE6_STATE_SIZE 256
#define DEC_DENSE7_OUT_SIZE 256
#define DEC_DENSE8_OUT_SIZE 256
#define DEC_FINAL_OUT_SIZE 80
#define DRED_DEC_MAX_RNN_NEURONS 256

struct RDOVAEDec {
  DenseLayer state1, state2, state3;
  DenseLayer dec_dense1, dec_dense3, dec_dense5, dec_dense7, dec_dense8, dec_final;
  GRULayer dec_dense2, dec_dense4, dec_dense6;
};

int init_rdovaedec(RDOVAEDec *model, const WeightArray *arrays);

#endif
