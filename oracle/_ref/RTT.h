#pragma once
#include <stdint.h>

/* Inferred minimum I/O channel counts (from splN usage / pin declarations) */
#define DSPJSFX_INPUT_CHANNELS 8
#define DSPJSFX_OUTPUT_CHANNELS 2
#define DSPJSFX_PROCESS_CHANNELS 8
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
    double vars[166];
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
#define DSPJSFX_VARS_COUNT 166
static const DSPJSFX_VarDesc DSPJSFX_VARS[166] = {
    {"Eo", 0},
    {"Eo_db", 1},
    {"Eo_fast", 2},
    {"Eo_slow", 3},
    {"Ev", 4},
    {"Ev_fast", 5},
    {"Ev_slow", 6},
    {"Ev_slow_db", 7},
    {"Ex_raw", 8},
    {"Ex_use", 9},
    {"Ey", 10},
    {"age_delay_ms", 11},
    {"age_delay_samps", 12},
    {"age_ms", 13},
    {"age_samps", 14},
    {"age_samps_max", 15},
    {"age_strength_db", 16},
    {"age_w", 17},
    {"age_wet_on_lin", 18},
    {"atk_a", 19},
    {"atk_ms", 20},
    {"eps", 21},
    {"floor_db", 22},
    {"floor_lin", 23},
    {"g", 24},
    {"gr_db", 25},
    {"gr_per_db", 26},
    {"grace_ms", 27},
    {"grace_samps", 28},
    {"guard_atk_a", 29},
    {"guard_atk_ms", 30},
    {"guard_rel_a", 31},
    {"guard_rel_ms", 32},
    {"guard_tgt", 33},
    {"half", 34},
    {"hold_ms", 35},
    {"hold_samps", 36},
    {"hold_samps_set", 37},
    {"knee_db", 38},
    {"maxduck_db", 39},
    {"mem_atk_a", 40},
    {"mem_atk_ms", 41},
    {"mix", 42},
    {"oL", 43},
    {"oR", 44},
    {"o_active", 45},
    {"o_credit", 46},
    {"o_env2", 47},
    {"o_exc", 48},
    {"o_exc_db", 49},
    {"o_exc_knee_db", 50},
    {"o_exc_norm", 51},
    {"o_exc_over", 52},
    {"o_exc_perm", 53},
    {"o_exc_thr_db", 54},
    {"o_fast2", 55},
    {"o_fast_a", 56},
    {"o_fast_ms", 57},
    {"o_half", 58},
    {"o_lvl_off_db", 59},
    {"o_lvl_on_db", 60},
    {"o_lvl_perm", 61},
    {"o_lvl_t", 62},
    {"o_mem", 63},
    {"o_mem_ms", 64},
    {"o_mem_rel_a", 65},
    {"o_on_db", 66},
    {"o_on_lin", 67},
    {"o_p", 68},
    {"o_perm", 69},
    {"o_perm_tgt", 70},
    {"o_ref", 71},
    {"o_slow2", 72},
    {"o_slow_a", 73},
    {"o_slow_ms", 74},
    {"o_t", 75},
    {"off_db", 76},
    {"off_lin", 77},
    {"operm_atk_a", 78},
    {"operm_atk_ms", 79},
    {"operm_rel_a", 80},
    {"operm_rel_ms", 81},
    {"other_auth", 82},
    {"other_auth_eff", 83},
    {"other_suppress", 84},
    {"over", 85},
    {"over_eff", 86},
    {"prev_ref_off", 87},
    {"procL", 88},
    {"procR", 89},
    {"ratio", 90},
    {"rdB", 91},
    {"ref_off", 92},
    {"rel_a", 93},
    {"rel_in_a", 94},
    {"rel_in_ms", 95},
    {"rel_ms", 96},
    {"rms_a", 97},
    {"rms_ms", 98},
    {"sens", 99},
    {"t", 100},
    {"tail_atk_a", 101},
    {"tail_atk_ms", 102},
    {"tail_candidate", 103},
    {"tail_floor_strength", 104},
    {"tail_live_ref", 105},
    {"tail_match_ms", 106},
    {"tail_nat", 107},
    {"tail_ref", 108},
    {"tail_ref_a", 109},
    {"tail_samps", 110},
    {"tail_seed_from_wet", 111},
    {"tail_seed_ratio_db", 112},
    {"tail_tau_ms", 113},
    {"tail_thr_bonus_db", 114},
    {"tail_w", 115},
    {"tgt_a", 116},
    {"tgt_db", 117},
    {"tgt_db_sm", 118},
    {"tgt_ms", 119},
    {"thr_db", 120},
    {"thr_eff_db", 121},
    {"vL", 122},
    {"vR", 123},
    {"v_active", 124},
    {"v_credit", 125},
    {"v_env2", 126},
    {"v_exc", 127},
    {"v_exc_db", 128},
    {"v_exc_knee_db", 129},
    {"v_exc_norm", 130},
    {"v_exc_over", 131},
    {"v_exc_perm", 132},
    {"v_exc_thr_db", 133},
    {"v_fast2", 134},
    {"v_fast_a", 135},
    {"v_fast_ms", 136},
    {"v_guard", 137},
    {"v_half", 138},
    {"v_mem", 139},
    {"v_mem_ms", 140},
    {"v_mem_rel_a", 141},
    {"v_on_db", 142},
    {"v_on_lin", 143},
    {"v_p", 144},
    {"v_perm", 145},
    {"v_perm_tgt", 146},
    {"v_ref", 147},
    {"v_slow2", 148},
    {"v_slow_a", 149},
    {"v_slow_ms", 150},
    {"v_sus", 151},
    {"v_sus_perm", 152},
    {"v_sus_t", 153},
    {"v_sustain_floor", 154},
    {"v_sustain_off_db", 155},
    {"v_sustain_on_db", 156},
    {"v_t", 157},
    {"vperm_atk_a", 158},
    {"vperm_atk_ms", 159},
    {"vperm_rel_a", 160},
    {"vperm_rel_ms", 161},
    {"wetL", 162},
    {"wetR", 163},
    {"wet_env2", 164},
    {"wet_p", 165},
};

/* @gfx user-var sync flags (vars[] index).
   bit0: sync audio-owned updates into the @gfx VM (audio -> gfx)
   bit1: feed @gfx-authored writes back into the DSP state (gfx -> audio)
   Mode: legacy */
#define DSPJSFX_GFX_VAR_FLAG_TO_GFX 1u
#define DSPJSFX_GFX_VAR_FLAG_FROM_GFX 2u
#define DSPJSFX_GFX_VAR_FLAGS_COUNT 166
static const uint8_t DSPJSFX_GFX_VAR_FLAGS[166] = {
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
