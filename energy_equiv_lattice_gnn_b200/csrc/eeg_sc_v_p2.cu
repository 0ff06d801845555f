// symmetric contraction, variant p2: 2 node(s) per thread, packed FFMA2, 128 threads per channel, 4 channels per CTA (see eeg_sc_impl.cuh)
#define EEG_SC_V_NAME p2
#define EEG_SC_V_NPT 2
#define EEG_SC_V_PACKED 1
#define EEG_SC_V_TPC 128
#define EEG_SC_V_CH 4
#include "eeg_sc_impl.cuh"
