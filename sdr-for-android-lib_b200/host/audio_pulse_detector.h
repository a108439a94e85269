// host/audio_pulse_detector.h — the part of the reference's AudioPulseDetector interface that crosses the
// SSBProcessor boundary (src/ssb/audio_pulse_detector.h:16-38): its Config, field for field with the same
// defaults, so that SSBProcessor::setPulseConfig(const AudioPulseDetector::Config&) keeps its signature.
// The detector itself runs on the GPU as SDRB_STAGE_AUDIO_PULSE (csrc/pulse.cu).
#pragma once

class AudioPulseDetector {
public:
    enum class PulseLevel { NONE = 0, LOW = 1, MEDIUM = 2, STRONG = 3 };

    struct Config {
        float sampleRate    = 48000.f;
        float fMin          = 1500.f;
        float fMax          = 4000.f;
        float fsEnergy      = 100.f;
        float smoothCutoff  = 5.f;
        float zDefaultS     = 0.666f;
        float tTargetInit   = 1.75f;
        float dtTolS        = 0.150f;
        float snrMin        = 1.0f;
        float snrRhythm     = 1.1f;
        float snrStrong     = 2.0f;
        float dispersionMax = 1.3f;
        int   sumNMax       = 7;
        float liveWindowT   = 4.0f;
        float liveDivisor   = 3.0f;
        int   noiseRefFar   = 80;
        int   noiseRefNear  = 40;
    };
};
