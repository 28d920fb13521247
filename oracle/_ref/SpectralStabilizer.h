#pragma once
#include <stdint.h>

/* Inferred minimum I/O channel counts (from splN usage / pin declarations) */
#define DSPJSFX_INPUT_CHANNELS 2
#define DSPJSFX_OUTPUT_CHANNELS 2
#define DSPJSFX_PROCESS_CHANNELS 2
#define DSPJSFX_USES_MIDI 0
#define DSPJSFX_USES_GMEM 0
#define DSPJSFX_USES_MSG 0
#define DSPJSFX_USES_MSG_BUFFERS 0
#define DSPJSFX_USES_SAMPLE_POOL 0
#define DSPJSFX_USES_RAW_SAMPLE_READ 0
#define DSPJSFX_USES_SAMPLE_EXPORT_MEM 0
#define DSPJSFX_USES_LEGACY_FILE_IO 0
#define DSPJSFX_COMM_API_VERSION 1
#define DSPJSFX_SAMPLE_POOL_API_VERSION 1
#define DSPJSFX_ACCEPTS_MIDI_INPUT 0
#define DSPJSFX_PRODUCES_MIDI_OUTPUT 0
#define DSPJSFX_HAS_SAMPLE_SECTION 1
#define DSPJSFX_PLUGIN_KIND "audio_effect"


#ifdef __cplusplus
extern "C" {
#endif

typedef struct DSPJSFX_MidiEvent {
    int32_t sampleOffset;
    int32_t msg1;
    int32_t msg2;
    int32_t msg3;
} DSPJSFX_MidiEvent;

typedef struct DSPJSFX_State {
    double spl[64];
    double sliders[64];
    double vars[108];
    double* mem;
    int64_t memN;
    double srate;
    double samplesblock;
    DSPJSFX_MidiEvent* midiIn;
    int32_t midiInCount;
    int32_t midiInReadIndex;
    int32_t midiInCapacity;
    DSPJSFX_MidiEvent* midiOut;
    int32_t midiOutCount;
    int32_t midiOutCapacity;
    int32_t currentBlockSize;
    double currentSampleRate;
    int32_t pendingNoteCleanup;
    int32_t midiInDropped;
    int32_t midiOutDropped;
    int32_t midiInCountLastBlock;
    int32_t midiOutCountLastBlock;
    int32_t midiInPeak;
    int32_t midiOutPeak;
    int64_t pendingSliderChangeMask;
    int64_t pendingSliderAutomateMask;
    int64_t pendingSliderAutomateEndMask;
    uint32_t randMT[624];
    uint32_t randIndex;
    int64_t sliderVisibleMask;
    int32_t sliderVisibilityInit;
    void* runtimeOpaque;
    double midi_bus;
    double ext_midi_bus;
} DSPJSFX_State;

/* User vars (name -> vars[] index) */
typedef struct DSPJSFX_VarDesc { const char* name; int32_t index; } DSPJSFX_VarDesc;
#define DSPJSFX_VARS_COUNT 108
static const DSPJSFX_VarDesc DSPJSFX_VARS[108] = {
    {"__fnlocal__init__biquad_bp_coefs__a0", 0},
    {"__fnlocal__init__biquad_bp_coefs__a1", 1},
    {"__fnlocal__init__biquad_bp_coefs__a2", 2},
    {"__fnlocal__init__biquad_bp_coefs__alpha", 3},
    {"__fnlocal__init__biquad_bp_coefs__b0", 4},
    {"__fnlocal__init__biquad_bp_coefs__b1", 5},
    {"__fnlocal__init__biquad_bp_coefs__b2", 6},
    {"__fnlocal__init__biquad_bp_coefs__w0", 7},
    {"__fnlocal__init__erb_cf__cf", 8},
    {"__fnlocal__init__erb_cf__ear_max", 9},
    {"__fnlocal__init__erb_cf__ear_min", 10},
    {"__fnlocal__init__erb_cf__t", 11},
    {"__fnlocal__init__fill_gauss_weights__j", 12},
    {"__fnlocal__init__fill_gauss_weights__k", 13},
    {"__fnlocal__init__fill_gauss_weights__sum", 14},
    {"__fnlocal__init__fill_gauss_weights__w", 15},
    {"__fnlocal__sample__biquad_proc__a1", 16},
    {"__fnlocal__sample__biquad_proc__a2", 17},
    {"__fnlocal__sample__biquad_proc__b0", 18},
    {"__fnlocal__sample__biquad_proc__b1", 19},
    {"__fnlocal__sample__biquad_proc__b2", 20},
    {"__fnlocal__sample__biquad_proc__y", 21},
    {"__fnlocal__sample__biquad_proc__z1", 22},
    {"__fnlocal__sample__biquad_proc__z2", 23},
    {"__fnlocal__slider__fill_gauss_weights__j", 24},
    {"__fnlocal__slider__fill_gauss_weights__k", 25},
    {"__fnlocal__slider__fill_gauss_weights__sum", 26},
    {"__fnlocal__slider__fill_gauss_weights__w", 27},
    {"acc_n", 28},
    {"acc_w", 29},
    {"action_amt", 30},
    {"atk_a", 31},
    {"atk_ms", 32},
    {"att_db", 33},
    {"bands", 34},
    {"bi", 35},
    {"bj", 36},
    {"bw", 37},
    {"cf", 38},
    {"d1", 39},
    {"d2", 40},
    {"dbfac", 41},
    {"depth", 42},
    {"e", 43},
    {"elog", 44},
    {"env", 45},
    {"env_a", 46},
    {"env_ms", 47},
    {"g_apply", 48},
    {"g_next", 49},
    {"g_prev", 50},
    {"g_t", 51},
    {"gauss_radius", 52},
    {"idx_b", 53},
    {"inL", 54},
    {"inR", 55},
    {"k", 56},
    {"k_ratio", 57},
    {"kneeadj", 58},
    {"ln10", 59},
    {"low_min_gain", 60},
    {"m", 61},
    {"m_inst", 62},
    {"maxHz", 63},
    {"maxHz_eff", 64},
    {"max_att_db", 65},
    {"mb", 66},
    {"mgate", 67},
    {"minHz", 68},
    {"min_gain_lin", 69},
    {"mot_a", 70},
    {"mot_ms", 71},
    {"ofs_bq", 72},
    {"ofs_bw", 73},
    {"ofs_cf", 74},
    {"ofs_d1", 75},
    {"ofs_d2", 76},
    {"ofs_envlin", 77},
    {"ofs_envlog", 78},
    {"ofs_gn", 79},
    {"ofs_gs", 80},
    {"ofs_gt", 81},
    {"ofs_gw", 82},
    {"ofs_mot", 83},
    {"ofs_q", 84},
    {"ofs_res", 85},
    {"ofs_z1L", 86},
    {"ofs_z1R", 87},
    {"ofs_z2L", 88},
    {"ofs_z2R", 89},
    {"onset", 90},
    {"onset_thr", 91},
    {"outL", 92},
    {"outR", 93},
    {"prev_d1", 94},
    {"prev_elog", 95},
    {"q", 96},
    {"r", 97},
    {"rel_a", 98},
    {"rel_ms", 99},
    {"res2db", 100},
    {"res_clip", 101},
    {"rk", 102},
    {"two_pi", 103},
    {"wn", 104},
    {"ww", 105},
    {"yL", 106},
    {"yR", 107},
};

/* @gfx user-var sync flags (vars[] index).
   bit0: sync audio-owned updates into the @gfx VM (audio -> gfx)
   bit1: feed @gfx-authored writes back into the DSP state (gfx -> audio)
   Mode: legacy */
#define DSPJSFX_GFX_VAR_FLAG_TO_GFX 1u
#define DSPJSFX_GFX_VAR_FLAG_FROM_GFX 2u
#define DSPJSFX_GFX_VAR_FLAGS_COUNT 108
static const uint8_t DSPJSFX_GFX_VAR_FLAGS[108] = {
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
};

/* String literals (opaque runtime handles used by MIDI/string helpers) */
typedef struct DSPJSFX_StringLiteralDesc { int64_t handle; int32_t length; const uint8_t* data; } DSPJSFX_StringLiteralDesc;
#define DSPJSFX_STRING_LITERALS_COUNT 0
static const uint8_t DSPJSFX_STRING_LITERAL_BYTES_0[1] = { 0 };
static const DSPJSFX_StringLiteralDesc DSPJSFX_STRING_LITERALS[1] = {
    { 0LL, 0, DSPJSFX_STRING_LITERAL_BYTES_0 },
};

/* Sections */
void jsfx_init(DSPJSFX_State* st);
void jsfx_slider(DSPJSFX_State* st);
void jsfx_block(DSPJSFX_State* st);
void jsfx_sample(DSPJSFX_State* st);

/* Entry point intended to be called from JUCE processBlock().
   inputs/outputs are arrays of channel pointers (non-interleaved). */
void jsfx_process_block(DSPJSFX_State* st,
                        const float* const* inputs,
                        float* const* outputs,
                        int32_t numChannels,
                        int32_t numSamples);

/* Runtime hook required by mem[] growth checks. You must provide this when linking.
   Even if you never exceed memN, the symbol must exist. */
void jsfx_ensure_mem(DSPJSFX_State* st, int64_t needed);
int jsfx_midirecv(DSPJSFX_State* st, double* offset, double* msg1, double* msg2, double* msg3);
int jsfx_midirecv_msg23(DSPJSFX_State* st, double* offset, double* msg1, double* msg23);
int jsfx_midirecv_buf(DSPJSFX_State* st, double* offset, double buf, double maxlen);
int jsfx_midirecv_str(DSPJSFX_State* st, double* offset, double* strSlot);
int jsfx_midisend(DSPJSFX_State* st, double offset, double msg1, double msg2, double msg3);
int jsfx_midisend_msg23(DSPJSFX_State* st, double offset, double msg1, double msg23);
int jsfx_midisend_buf(DSPJSFX_State* st, double offset, double buf, double len);
int jsfx_midisend_str(DSPJSFX_State* st, double offset, double strHandle);
int jsfx_midisyx(DSPJSFX_State* st, double offset, double msgptr, double len);
int jsfx_strlen(DSPJSFX_State* st, double strHandle);
int jsfx_str_getchar(DSPJSFX_State* st, double strHandle, double index);
int jsfx_sliderchange(DSPJSFX_State* st, double sliderMask);
int jsfx_slider_automate(DSPJSFX_State* st, double sliderMask, double endTouch);
double jsfx_slider_next_chg(DSPJSFX_State* st, double sliderIndex, double* outValue);
double jsfx_instance_id(DSPJSFX_State* st);
int jsfx_instance_uid(DSPJSFX_State* st, double* outStr);
int jsfx_instance_set_name(DSPJSFX_State* st, double strHandle);
int jsfx_instance_get_name(DSPJSFX_State* st, double* outStr);
int jsfx_track_name(DSPJSFX_State* st, double* outStr);
double jsfx_track_name_available(DSPJSFX_State* st);
double jsfx_track_name_seq(DSPJSFX_State* st);
int jsfx_comm_join(DSPJSFX_State* st, double domainHandle);
int jsfx_gmem_attach(DSPJSFX_State* st, double nameHandle);
int jsfx_gmem_attach_size(DSPJSFX_State* st, double nameHandle, double cells);
double jsfx_gmem_size(DSPJSFX_State* st);
double jsfx_gmem_load(DSPJSFX_State* st, double idx);
double jsfx_gmem_store(DSPJSFX_State* st, double idx, double value);
int jsfx_gmem_get(DSPJSFX_State* st, double dstBase, double srcIdx, double count);
int jsfx_gmem_put(DSPJSFX_State* st, double dstIdx, double srcBase, double count);
int jsfx_gmem_fill(DSPJSFX_State* st, double dstIdx, double value, double count);
int jsfx_gmem_zero(DSPJSFX_State* st, double dstIdx, double count);
int jsfx_gmem_copy(DSPJSFX_State* st, double dstIdx, double srcIdx, double count);
double jsfx_gmem_seq(DSPJSFX_State* st, double page);
double jsfx_gmem_page(DSPJSFX_State* st, double idx);
int jsfx_msg_subscribe(DSPJSFX_State* st, double chanHandle);
int jsfx_msg_unsubscribe(DSPJSFX_State* st, double chanHandle);
int jsfx_msg_advertise(DSPJSFX_State* st, double chanHandle, double caps);
int jsfx_msg_send(DSPJSFX_State* st, double chanHandle, double tag, double a, double b, double c, double d);
int jsfx_msg_sendto(DSPJSFX_State* st, double targetId, double chanHandle, double tag, double a, double b, double c, double d);
double jsfx_msg_avail(DSPJSFX_State* st, double chanHandle);
double jsfx_msg_kind(DSPJSFX_State* st, double chanHandle);
int jsfx_msg_recv(DSPJSFX_State* st, double chanHandle, double* src, double* tag, double* a, double* b, double* c, double* d);
int jsfx_msg_send_buf(DSPJSFX_State* st, double chanHandle, double tag, double srcBase, double len);
int jsfx_msg_sendto_buf(DSPJSFX_State* st, double targetId, double chanHandle, double tag, double srcBase, double len);
int jsfx_msg_recv_buf(DSPJSFX_State* st, double chanHandle, double* src, double* tag, double dstBase, double maxLen);
double jsfx_msg_length(DSPJSFX_State* st);
double jsfx_msg_dropped(DSPJSFX_State* st, double chanHandle);
int jsfx_msg_clear(DSPJSFX_State* st, double chanHandle);
double jsfx_msg_peer_count(DSPJSFX_State* st, double chanHandle, double role);
double jsfx_msg_peer_id(DSPJSFX_State* st, double chanHandle, double role, double index);
int jsfx_msg_peer_name(DSPJSFX_State* st, double peerId, double* outStr);
int jsfx_msg_peer_uid(DSPJSFX_State* st, double peerId, double* outStr);
double jsfx_msg_peer_caps(DSPJSFX_State* st, double peerId);
double jsfx_msg_peer_alive(DSPJSFX_State* st, double peerId);
double jsfx_sample_pool_from_slot(DSPJSFX_State* st, double slot, double nameHandle);
double jsfx_sample_pool_set_mode(DSPJSFX_State* st, double pool, double mode);
double jsfx_sample_pool_set_budget_mb(DSPJSFX_State* st, double pool, double mb);
double jsfx_sample_pool_commit(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_state(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_selected(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_loaded(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_failed(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_ram_mb(DSPJSFX_State* st, double pool);
double jsfx_sample_pool_generation(DSPJSFX_State* st, double pool);
double jsfx_sample_get(DSPJSFX_State* st, double pool, double index);
double jsfx_sample_len(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_channels(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_srate(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_peak(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_rms(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_name(DSPJSFX_State* st, double pool, double sampleId, double* outStr);
double jsfx_sample_read(DSPJSFX_State* st, double pool, double sampleId, double channel, double frame);
double jsfx_sample_read_interp(DSPJSFX_State* st, double pool, double sampleId, double channel, double phase);
double jsfx_sample_read2(DSPJSFX_State* st, double pool, double sampleId, double phase, double* outL, double* outR);
double jsfx_sample_read2_interp(DSPJSFX_State* st, double pool, double sampleId, double phase, double* outL, double* outR);
double jsfx_sample_preview_bins(DSPJSFX_State* st, double pool, double sampleId);
double jsfx_sample_preview_read(DSPJSFX_State* st, double pool, double sampleId, double bin, double* minValue, double* maxValue, double* rmsValue);
double jsfx_sample_export_mem(DSPJSFX_State* st, double pool, double sampleId, double dstBase, double srcFrame, double frameCount);
double jsfx_sample_export_mem2(DSPJSFX_State* st, double pool, double sampleId, double dstBase, double srcFrame, double frameCount);

#ifdef __cplusplus
}
#endif
