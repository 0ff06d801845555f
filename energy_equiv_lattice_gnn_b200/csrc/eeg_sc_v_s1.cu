// symmetric contraction, variant s1: 1 node(s) per thread, scalar FFMA, 128 threads per channel, 4 channels per CTA (see eeg_sc_impl.cuh)
#define EEG_SC_V_NAME s1
#define EEG_SC_V_NPT 1
#define EEG_SC_V_PACKED 0
#define EEG_SC_V_TPC 128
#define EEG_SC_V_CH 4
#include "eeg_sc_impl.cuh"
