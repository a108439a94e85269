/* oracle/sdr_oracle_airspy.c — plain-C restatement of the Airspy real->IQ front end
 * (SURVEY.md 8(f) row 1).  TEST INFRASTRUCTURE ONLY.
 *
 *   12-bit real ADC samples in uint16 at twice the complex rate
 *     -> float      convert_samples_float   third_party/libairspy/libairspy/src/airspy.c:311-321
 *     -> DC removal remove_dc               .../iqconverter_float.c:418-431
 *     -> fs/4 shift translate_fs_4          .../iqconverter_float.c:433-469
 *     -> half-band  fir_interleaved_24      .../iqconverter_float.c:287-326 (47-tap kernel, filters.h:28-79)
 *     -> delay      delay_interleaved       .../iqconverter_float.c:391-414
 * Pinned bit-exactly against the reference's own iqconverter_float.c compiled into
 * oracle/_ref/libsdr_ref.so (tests/test_oracle_airspy.py). */
#include <stdint.h>
#include <string.h>

#define HB_LEN 47
#define FIR_LEN 24   /* len/2 + 1, iqconverter_float.c:88 */

/* the 24 non-zero even taps of HB_KERNEL_FLOAT (filters.h:30-78); only the first 12 are used (symmetry) */
static const float kHb[12] = {
    -0.000998606272947510f, 0.001695637278417295f, -0.003054430179754289f, 0.005055504379767936f,
    -0.007901319195893647f, 0.011873357051047719f, -0.017411159379930066f, 0.025304817427568772f,
    -0.037225225204559217f, 0.057533286997004301f, -0.102327462004259350f, 0.317034472508947400f};

typedef struct {
    float avg;            /* cnv->avg */
    float even[FIR_LEN];  /* last 24 translated even samples, even[0] = newest */
    float odd[12];        /* delay line contents, odd[0] = newest */
} sdr_oracle_airspy_state;

void sdr_oracle_airspy_init(sdr_oracle_airspy_state* s) { memset(s, 0, sizeof *s); }

/* raw: n_real uint16 samples (n_real % 4 == 0); iq_out: n_real/2 complex floats (interleaved) */
void sdr_oracle_airspy_process(sdr_oracle_airspy_state* s, const uint16_t* raw, int n_real, float* iq_out) {
    const float scale = 1.0f / 2048.0f;   /* SAMPLE_SCALE, airspy.c:58-62 */
    const float hbc = 0.5f;               /* hb_kernel[len/2], iqconverter_float.c:89 */
    float avg = s->avg;
    for (int i = 0; i < n_real; i++) {
        float x = (float)((int)raw[i] - 2048) * scale;   /* airspy.c:316 */
        x -= avg;                                        /* iqconverter_float.c:425-426 */
        avg += 0.01f * x;
        switch (i & 3) {                                 /* iqconverter_float.c:459-464 */
            case 0: x = -x; break;
            case 1: x = -x * hbc; break;
            case 2: break;
            default: x = x * hbc; break;
        }
        if ((i & 1) == 0) {
            memmove(s->even + 1, s->even, sizeof(float) * (FIR_LEN - 1));
            s->even[0] = x;
            /* fir_interleaved_24: left-to-right sum of twelve products (iqconverter_float.c:303-314) */
            float acc = kHb[0] * (s->even[0] + s->even[23]);
            for (int k = 1; k < 12; k++) acc = acc + kHb[k] * (s->even[k] + s->even[23 - k]);
            iq_out[i] = acc;
        } else {
            const float res = s->odd[11];                /* delay of len>>1 = 12 (iqconverter_float.c:399-412) */
            memmove(s->odd + 1, s->odd, sizeof(float) * 11);
            s->odd[0] = x;
            iq_out[i] = res;
        }
    }
    s->avg = avg;
}
