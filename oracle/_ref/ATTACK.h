#pragma once
#include <stdint.h>

/* Inferred minimum I/O channel counts (from splN usage / pin declarations) */
#define DSPJSFX_INPUT_CHANNELS 2
#define DSPJSFX_OUTPUT_CHANNELS 4
#define DSPJSFX_PROCESS_CHANNELS 4
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
    double vars[238];
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
#define DSPJSFX_VARS_COUNT 238
static const DSPJSFX_VarDesc DSPJSFX_VARS[238] = {
    {"Q", 0},
    {"a1a", 1},
    {"a1b", 2},
    {"a2a", 3},
    {"a2b", 4},
    {"aL0", 5},
    {"aL1", 6},
    {"aL2", 7},
    {"aPrevL0", 8},
    {"aPrevL1", 9},
    {"aPrevL2", 10},
    {"aPrevR0", 11},
    {"aPrevR1", 12},
    {"aPrevR2", 13},
    {"aR0", 14},
    {"aR1", 15},
    {"aR2", 16},
    {"aa0", 17},
    {"aa1", 18},
    {"aa2", 19},
    {"alpha", 20},
    {"b0a", 21},
    {"b0b", 22},
    {"b1a", 23},
    {"b1b", 24},
    {"b2a", 25},
    {"b2b", 26},
    {"bb0", 27},
    {"bb1", 28},
    {"bb2", 29},
    {"body_db", 30},
    {"brutal_bonus", 31},
    {"cap", 32},
    {"crack_db", 33},
    {"cw", 34},
    {"db2lin", 35},
    {"db_high", 36},
    {"db_low", 37},
    {"db_mid", 38},
    {"delta_boost", 39},
    {"edge_db", 40},
    {"envG_body", 41},
    {"envG_crack", 42},
    {"envG_edge", 43},
    {"envG_punch", 44},
    {"envS_body", 45},
    {"envS_crack_high", 46},
    {"envS_crack_mid", 47},
    {"envS_edge_high", 48},
    {"envS_edge_mid", 49},
    {"envS_punch_low", 50},
    {"envS_punch_mid", 51},
    {"env_body_used", 52},
    {"env_crack_high", 53},
    {"env_crack_mid", 54},
    {"env_edge_high", 55},
    {"env_edge_mid", 56},
    {"env_punch_low", 57},
    {"env_punch_mid", 58},
    {"eps", 59},
    {"fc_low", 60},
    {"fc_mid", 61},
    {"flavor", 62},
    {"gB", 63},
    {"gC", 64},
    {"gCh", 65},
    {"gCm", 66},
    {"gE", 67},
    {"gEh", 68},
    {"gEm", 69},
    {"gH", 70},
    {"gL", 71},
    {"gM", 72},
    {"gP", 73},
    {"gPl", 74},
    {"gPm", 75},
    {"graw", 76},
    {"highL", 77},
    {"highR", 78},
    {"jL0", 79},
    {"jL1", 80},
    {"jL2", 81},
    {"jR0", 82},
    {"jR1", 83},
    {"jR2", 84},
    {"k_body", 85},
    {"k_crack", 86},
    {"k_edge", 87},
    {"k_punch", 88},
    {"lowL", 89},
    {"lowR", 90},
    {"lp250L", 91},
    {"lp250R", 92},
    {"lp4kL", 93},
    {"lp4kR", 94},
    {"m", 95},
    {"mag0_0", 96},
    {"mag0_1", 97},
    {"mag0_2", 98},
    {"mag1_0", 99},
    {"mag1_1", 100},
    {"mag1_2", 101},
    {"mag2_0", 102},
    {"mag2_1", 103},
    {"mag2_2", 104},
    {"mag3_0", 105},
    {"mag3_1", 106},
    {"mag3_2", 107},
    {"magG_body", 108},
    {"magG_crack", 109},
    {"magG_edge", 110},
    {"magG_punch", 111},
    {"max_body_db", 112},
    {"max_crack_db", 113},
    {"max_edge_db", 114},
    {"max_punch_db", 115},
    {"midL", 116},
    {"midR", 117},
    {"oldvS", 118},
    {"pi", 119},
    {"punch_db", 120},
    {"r0_0", 121},
    {"r0_1", 122},
    {"r0_2", 123},
    {"r1_0", 124},
    {"r1_1", 125},
    {"r1_2", 126},
    {"r2_0", 127},
    {"r2_1", 128},
    {"r2_2", 129},
    {"r3_0", 130},
    {"r3_1", 131},
    {"r3_2", 132},
    {"ref0_0", 133},
    {"ref0_1", 134},
    {"ref0_2", 135},
    {"ref0_base", 136},
    {"ref1_0", 137},
    {"ref1_1", 138},
    {"ref1_2", 139},
    {"ref1_base", 140},
    {"ref2_0", 141},
    {"ref2_1", 142},
    {"ref2_2", 143},
    {"ref2_base", 144},
    {"ref3_0", 145},
    {"ref3_1", 146},
    {"ref3_2", 147},
    {"ref3_base", 148},
    {"refG_body", 149},
    {"refG_crack", 150},
    {"refG_edge", 151},
    {"refG_punch", 152},
    {"refc0_0", 153},
    {"refc0_1", 154},
    {"refc0_2", 155},
    {"refc1_0", 156},
    {"refc1_1", 157},
    {"refc1_2", 158},
    {"refc2_0", 159},
    {"refc2_1", 160},
    {"refc2_2", 161},
    {"refc3_0", 162},
    {"refc3_1", 163},
    {"refc3_2", 164},
    {"refcG_body", 165},
    {"refcG_crack", 166},
    {"refcG_edge", 167},
    {"refcG_punch", 168},
    {"rel_body", 169},
    {"rel_body_ms", 170},
    {"rel_crack", 171},
    {"rel_crack_ms", 172},
    {"rel_edge", 173},
    {"rel_edge_ms", 174},
    {"rel_punch", 175},
    {"rel_punch_ms", 176},
    {"rg_body", 177},
    {"rg_crack", 178},
    {"rg_edge", 179},
    {"rg_punch", 180},
    {"rs_body", 181},
    {"s", 182},
    {"sPrevL0", 183},
    {"sPrevL1", 184},
    {"sPrevL2", 185},
    {"sPrevR0", 186},
    {"sPrevR1", 187},
    {"sPrevR2", 188},
    {"shp", 189},
    {"smooth_ms", 190},
    {"sw", 191},
    {"tPow", 192},
    {"tauH", 193},
    {"tauL", 194},
    {"tauM", 195},
    {"tau_mul_high", 196},
    {"tau_mul_low", 197},
    {"tau_mul_mid", 198},
    {"thr_body", 199},
    {"thr_body_db", 200},
    {"thr_crack", 201},
    {"thr_crack_db", 202},
    {"thr_crack_high_mul", 203},
    {"thr_crack_mid_mul", 204},
    {"thr_edge", 205},
    {"thr_edge_db", 206},
    {"thr_edge_high_mul", 207},
    {"thr_punch", 208},
    {"thr_punch_db", 209},
    {"tight", 210},
    {"tight_adj", 211},
    {"v", 212},
    {"vSL0", 213},
    {"vSL1", 214},
    {"vSL2", 215},
    {"vSR0", 216},
    {"vSR1", 217},
    {"vSR2", 218},
    {"vcoef0", 219},
    {"vcoef1", 220},
    {"vcoef2", 221},
    {"w", 222},
    {"w_crack_mid", 223},
    {"w_edge_high", 224},
    {"w_punch_mid", 225},
    {"xL", 226},
    {"xR", 227},
    {"yL", 228},
    {"yR", 229},
    {"z1aL", 230},
    {"z1aR", 231},
    {"z1bL", 232},
    {"z1bR", 233},
    {"z2aL", 234},
    {"z2aR", 235},
    {"z2bL", 236},
    {"z2bR", 237},
};

/* @gfx user-var sync flags (vars[] index).
   bit0: sync audio-owned updates into the @gfx VM (audio -> gfx)
   bit1: feed @gfx-authored writes back into the DSP state (gfx -> audio)
   Mode: legacy */
#define DSPJSFX_GFX_VAR_FLAG_TO_GFX 1u
#define DSPJSFX_GFX_VAR_FLAG_FROM_GFX 2u
#define DSPJSFX_GFX_VAR_FLAGS_COUNT 238
static const uint8_t DSPJSFX_GFX_VAR_FLAGS[238] = {
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
