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
    double vars[245];
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
#define DSPJSFX_VARS_COUNT 245
static const DSPJSFX_VarDesc DSPJSFX_VARS[245] = {
    {"AUDIO_AIR_ADD", 0},
    {"AUDIO_BODY_LOW", 1},
    {"AUDIO_BODY_MID", 2},
    {"AUDIO_BODY_OUT", 3},
    {"AUDIO_MOTION_BASE", 4},
    {"AUDIO_PRESENCE_ADD", 5},
    {"DEF_BODY_PCT", 6},
    {"DEF_CEILING_DB", 7},
    {"DEF_DISPLAY_RANGE", 8},
    {"DEF_DRIVE_DB", 9},
    {"DEF_GUARD", 10},
    {"DEF_LOOKAHEAD_MS", 11},
    {"DEF_MOTION_PCT", 12},
    {"DEF_PRESERVE_PCT", 13},
    {"DEF_RELEASE_MS", 14},
    {"DEF_RELIEF_PCT", 15},
    {"DEF_STEREO_LINK", 16},
    {"DEF_STYLE", 17},
    {"DEF_TAME_PCT", 18},
    {"HIST_SIZE", 19},
    {"MAX_LOOK", 20},
    {"RELIEF_ADD_DUCK", 21},
    {"RELIEF_BODY_DUCK", 22},
    {"RELIEF_BODY_VALLEY_DUCK", 23},
    {"RELIEF_FULL_GR_DB", 24},
    {"RELIEF_MIN_ADD_KEEP", 25},
    {"RELIEF_MIN_BODY_KEEP", 26},
    {"RELIEF_RELEASE_HOLD", 27},
    {"RELIEF_START_GR_DB", 28},
    {"RELIEF_TRANSIENT_GAIN", 29},
    {"RELIEF_VALLEY_DUCK", 30},
    {"STYLE_BODY_BODY", 31},
    {"STYLE_BODY_PRES", 32},
    {"STYLE_BODY_REL", 33},
    {"STYLE_BODY_TILT", 34},
    {"STYLE_CLEAN_BODY", 35},
    {"STYLE_CLEAN_PRES", 36},
    {"STYLE_CLEAN_REL", 37},
    {"STYLE_CLEAN_TILT", 38},
    {"STYLE_DENSE_BODY", 39},
    {"STYLE_DENSE_PRES", 40},
    {"STYLE_DENSE_REL", 41},
    {"STYLE_DENSE_TILT", 42},
    {"STYLE_PUNCH_BODY", 43},
    {"STYLE_PUNCH_PRES", 44},
    {"STYLE_PUNCH_REL", 45},
    {"STYLE_PUNCH_TILT", 46},
    {"STYLE_WILD_BODY", 47},
    {"STYLE_WILD_PRES", 48},
    {"STYLE_WILD_REL", 49},
    {"STYLE_WILD_TILT", 50},
    {"TAME_AIR_DEPTH", 51},
    {"TAME_BODY_MID_DEPTH", 52},
    {"TAME_FULL_GR_DB", 53},
    {"TAME_LOW_DEPTH", 54},
    {"TAME_MIN_ADD_KEEP", 55},
    {"TAME_PRESENCE_DEPTH", 56},
    {"TAME_START_GR_DB", 57},
    {"UI_BOTTOM_H", 58},
    {"UI_CEIL_BADGE_DB_PER_PX", 59},
    {"UI_DBL_FRAMES", 60},
    {"UI_DRIVE_BADGE_DB_PER_PX", 61},
    {"UI_TILE_GAP", 62},
    {"UI_TILE_H", 63},
    {"UI_TOP_H", 64},
    {"UI_TT_BG_ALPHA", 65},
    {"UI_TT_BLUE_LINE_H", 66},
    {"UI_TT_BODY_SIZE", 67},
    {"UI_TT_BOTTOM_EXTRA", 68},
    {"UI_TT_LINE_HEIGHT", 69},
    {"UI_TT_MIN_WIDTH", 70},
    {"UI_TT_OFFSET_X", 71},
    {"UI_TT_OFFSET_Y", 72},
    {"UI_TT_PADDING", 73},
    {"UI_TT_RED_LINE_H", 74},
    {"UI_TT_SCREEN_MARGIN", 75},
    {"UI_TT_SHADOW_ALPHA", 76},
    {"UI_TT_TITLE_GAP", 77},
    {"UI_TT_TITLE_SIZE", 78},
    {"UI_TT_WIDTH", 79},
    {"__fnlocal__init__onepole_coeff__f", 80},
    {"__fnlocal__init__update_params__hist_hz", 81},
    {"__fnlocal__init__update_params__rel_eff", 82},
    {"__fnlocal__init__update_params__style_i", 83},
    {"__fnlocal__sample__softsat__k", 84},
    {"__fnlocal__slider__onepole_coeff__f", 85},
    {"__fnlocal__slider__update_params__hist_hz", 86},
    {"__fnlocal__slider__update_params__rel_eff", 87},
    {"__fnlocal__slider__update_params__style_i", 88},
    {"absL", 89},
    {"absR", 90},
    {"add_lp120L", 91},
    {"add_lp120R", 92},
    {"add_lp1600L", 93},
    {"add_lp1600R", 94},
    {"add_lp420L", 95},
    {"add_lp420R", 96},
    {"add_lp6000L", 97},
    {"add_lp6000R", 98},
    {"airBandL", 99},
    {"airBandR", 100},
    {"atkcoef", 101},
    {"attack_samples", 102},
    {"bodyAddL", 103},
    {"bodyAddR", 104},
    {"bodyBandL", 105},
    {"bodyBandR", 106},
    {"bodyDrive", 107},
    {"bodySrcL", 108},
    {"bodySrcR", 109},
    {"body_amt", 110},
    {"body_det", 111},
    {"body_pct", 112},
    {"bodymul", 113},
    {"bufL", 114},
    {"bufR", 115},
    {"c120", 116},
    {"c1600", 117},
    {"c3000", 118},
    {"c350", 119},
    {"c420", 120},
    {"c6000", 121},
    {"ceiling_db", 122},
    {"ceiling_lin", 123},
    {"cur_gr_db", 124},
    {"delL", 125},
    {"delR", 126},
    {"depth_now", 127},
    {"detL", 128},
    {"detR", 129},
    {"det_lp120", 130},
    {"det_lp3000", 131},
    {"det_lp350", 132},
    {"det_peak_linked", 133},
    {"det_side_lp350", 134},
    {"det_tilt", 135},
    {"display_range_db", 136},
    {"drag_mode", 137},
    {"drive_db", 138},
    {"drive_lin", 139},
    {"fastrelcoef", 140},
    {"gainL", 141},
    {"gainR", 142},
    {"gr_amt", 143},
    {"gr_lin", 144},
    {"gr_show_db", 145},
    {"guard", 146},
    {"high_det", 147},
    {"histGR", 148},
    {"histPeak", 149},
    {"histRMS", 150},
    {"hist_count", 151},
    {"hist_skip", 152},
    {"hist_wpos", 153},
    {"inL", 154},
    {"inR", 155},
    {"jump_gain", 156},
    {"jump_peak", 157},
    {"last_mouse_cap", 158},
    {"last_mouse_y", 159},
    {"link_amt", 160},
    {"look_samps", 161},
    {"lookahead_ms", 162},
    {"low_det", 163},
    {"mainL", 164},
    {"mainR", 165},
    {"main_gr_db_pos", 166},
    {"meter_peak", 167},
    {"meter_peak_decay", 168},
    {"meter_rms2", 169},
    {"meter_rms_coef", 170},
    {"mid", 171},
    {"motionL", 172},
    {"motionR", 173},
    {"motion_amt", 174},
    {"motion_coef", 175},
    {"motion_pct", 176},
    {"native_sliders_hidden", 177},
    {"outL", 178},
    {"outR", 179},
    {"out_abs", 180},
    {"out_peak", 181},
    {"peak", 182},
    {"peak_acc", 183},
    {"power", 184},
    {"pre_gr_db_pos", 185},
    {"presAddL", 186},
    {"presAddR", 187},
    {"presBandL", 188},
    {"presBandR", 189},
    {"pres_det", 190},
    {"preserve", 191},
    {"preserve_pct", 192},
    {"presmul", 193},
    {"prevInL", 194},
    {"prevInR", 195},
    {"prevOutL", 196},
    {"prevOutR", 197},
    {"psy_peak", 198},
    {"randA", 199},
    {"randB", 200},
    {"rand_state", 201},
    {"range_db", 202},
    {"rdpos", 203},
    {"rel_now", 204},
    {"relcoef", 205},
    {"release_ms", 206},
    {"relief_add_scale", 207},
    {"relief_amt", 208},
    {"relief_body_scale", 209},
    {"relief_fast", 210},
    {"relief_fast_decay", 211},
    {"relief_level", 212},
    {"relief_pct", 213},
    {"relief_power", 214},
    {"relief_power_pre", 215},
    {"relief_slow", 216},
    {"relief_slow_coef", 217},
    {"relief_trans_gain", 218},
    {"relief_transient", 219},
    {"relief_valley", 220},
    {"relief_zone", 221},
    {"relief_zone_pre", 222},
    {"relmul", 223},
    {"rms_acc", 224},
    {"safety_gain", 225},
    {"safety_show_db", 226},
    {"side", 227},
    {"side_det", 228},
    {"stereo_link_pct", 229},
    {"style", 230},
    {"tame_air_scale", 231},
    {"tame_amt", 232},
    {"tame_body_mid_scale", 233},
    {"tame_low_scale", 234},
    {"tame_pct", 235},
    {"tame_power", 236},
    {"tame_pres_scale", 237},
    {"tame_zone", 238},
    {"targetL", 239},
    {"targetL_unlinked", 240},
    {"targetR", 241},
    {"targetR_unlinked", 242},
    {"target_linked", 243},
    {"wrpos", 244},
};

/* @gfx user-var sync flags (vars[] index).
   bit0: sync audio-owned updates into the @gfx VM (audio -> gfx)
   bit1: feed @gfx-authored writes back into the DSP state (gfx -> audio)
   Mode: legacy */
#define DSPJSFX_GFX_VAR_FLAG_TO_GFX 1u
#define DSPJSFX_GFX_VAR_FLAG_FROM_GFX 2u
#define DSPJSFX_GFX_VAR_FLAGS_COUNT 245
static const uint8_t DSPJSFX_GFX_VAR_FLAGS[245] = {
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
    3,
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
