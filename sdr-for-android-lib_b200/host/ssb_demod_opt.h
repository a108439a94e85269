// host/ssb_demod_opt.h — header-compatible front end of the reference's SSB entry point
// (src/ssb/ssb_demod_opt.h:64-65) over the C-ABI of include/sdr_b200.h.
//
// processSSB_opt keeps ALL filter state in function statics in the reference
// (src/ssb/ssb_demod_opt.cpp:224-227,261,277), i.e. one stream per process; this front end
// keeps one process-wide GPU context (one stream, one reading per call) the same way.  The
// stage functions of the reference header (removeDC, iir2Process, ...) are internal steps of
// that chain; on the GPU they are fused into the SSB kernels and are not exported one by one.
#pragma once
#include <complex>
#include <cstdint>
#include <vector>

// iq is taken by value like the reference's signature; `pulse` is never written (the reference
// never writes it either).  Throws std::runtime_error on any CUDA failure.
void processSSB_opt(std::vector<std::complex<float>> iq, uint32_t sampleRate, bool upperSideband,
                    std::vector<int16_t>& pcmOut, bool& pulse, int mode);

// src/ssb/ssb_demod_opt.h:67 — the reference's implementation always answers false.
bool lastBlockContainsPulse();

// additive: CUDA device used by processSSB_opt's context (call before the first block), and a
// reset that forgets the carried filter state (the reference can only do that by restarting).
void sdrbSsbSetDevice(int device);
void sdrbSsbReset();

// additive: processSSB_opt followed by the AudioPulseDetector step the SSB worker runs on the PCM
// (src/ssb/ssb_processor.cpp:103-113), in one GPU call; *pulse_out receives the detector's state after
// this block.  new_cfg (may be NULL) rebuilds the detector first, like setPulseConfig.
#include "../../include/sdr_b200.h"
void sdrbSsbProcessWithPulse(std::vector<std::complex<float>> iq, uint32_t sampleRate, bool upperSideband,
                             std::vector<int16_t>& pcmOut, int mode, const sdrb_audio_pulse_config* new_cfg,
                             sdrb_pulse_result* pulse_out);
