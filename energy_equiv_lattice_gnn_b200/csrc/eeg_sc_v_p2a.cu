// symmetric contraction, variant p2a: packed FFMA2 node pairs, 128 threads per channel, 4 channels per CTA,
// cp.async-prefetched x tiles and direct global stores (see eeg_sc_impl.cuh)
#define EEG_SC_V_NAME p2a
#define EEG_SC_V_NPT 2
#define EEG_SC_V_PACKED 1
#define EEG_SC_V_TPC 128
#define EEG_SC_V_CH 4
#define EEG_SC_V_ASYNC 1
#include "eeg_sc_impl.cuh"
