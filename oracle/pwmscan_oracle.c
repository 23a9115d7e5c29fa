/*
 * TEST INFRASTRUCTURE ONLY -- plain-C restatement of the scoring loops of PWMScan's
 * pwm_scoring (explainn/pwmscan/pwm_scoring.c in the reference checkout), exposed as
 * library functions for the parity tests.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * Pinning: tests/test_oracle_golden.py compares these functions with the text output of
 * the reference tool itself (oracle/_ref/pwm_scoring, compiled from the reference source
 * by oracle/Makefile) stored in tests/golden/pwmscan_golden.json.
 *
 * Sequences are int8 codes A0 C1 G2 T3 other 4 (pwm_scoring.c:627-655); matrices have
 * 5 rows (row 4 = N: 0.25 for LPM :862-865, INT_MIN for integer PWM :877-879).
 */
#include <limits.h>
#include <stdint.h>

/* pwm_scoring.c:445-474 -- sum of window probabilities over one record */
double oracle_lpm_sum(const int8_t* seq, long long len, int w, const double* lpm /*[5][w]*/,
                      const double* bg /*[5]*/, int forward_only) {
    double sum = 0.0;
    for (long long i = 0; i <= len - w; i++) {
        double prod = 1.0, prod_rc = 1.0;
        for (int j = 0; j < w; j++) {
            int c = seq[i + j];
            prod = prod * lpm[c * w + j] / bg[c];
            if (!forward_only) {
                int r = (c == 4) ? 4 : 3 - c;
                prod_rc = prod_rc * lpm[r * w + (w - j - 1)] / bg[r];
            }
        }
        if (forward_only) sum = sum + prod;
        else sum = sum + prod + prod_rc;
    }
    return sum;
}

/* pwm_scoring.c:381-444 -- best single match; returns the number of later tied hits.
 * pos/ties hold the printed positions (start for '+', start + w for '-'). */
int oracle_lpm_best(const int8_t* seq, long long len, int w, const double* lpm, const double* bg,
                    int forward_only, double* best_out, int* pos_out, char* strand_out,
                    int* ties, int tie_cap) {
    double best = 0.0;
    int pos = 0, nties = 0;
    char strand = '+';
    for (long long i = 0; i <= len - w; i++) {
        double prod = 1.0, prod_rc = 1.0;
        for (int j = 0; j < w; j++) {
            int c = seq[i + j];
            prod = prod * lpm[c * w + j] / bg[c];
            if (!forward_only) {
                int r = (c == 4) ? 4 : 3 - c;
                prod_rc = prod_rc * lpm[r * w + (w - j - 1)] / bg[r];
            }
        }
        double max = forward_only ? prod : (prod > prod_rc ? prod : prod_rc);
        if (max > best) {
            best = max;
            pos = (int)i;
            nties = 0;
            if (!forward_only) {
                if (max == prod) strand = '+';
                else { strand = '-'; pos = (int)i + w; }
            }
        } else if (max == best && max != 0.0) {
            int p = (max == prod) ? (int)i : (int)i + w;
            if (nties < tie_cap) ties[nties] = p;
            nties++;
        }
    }
    *best_out = best; *pos_out = pos; *strand_out = strand;
    return nties;
}

/* pwm_scoring.c:477-573 -- integer PWM best hit (first maximum wins, '+' on strand ties);
 * unsigned arithmetic reproduces the two's-complement wrap of INT_MIN entries. */
void oracle_ipwm_best(const int8_t* seq, long long len, int w, const int32_t* pwm /*[5][w]*/,
                      int forward_only, int32_t* best_out, int32_t* pos_out, int8_t* strand_out) {
    int32_t best = INT_MIN, pos = 0;
    int8_t strand = 0;
    if (len < w) { *best_out = INT_MIN; *pos_out = -1; *strand_out = 0; return; }
    for (long long i = 0; i <= len - w; i++) {
        uint32_t score = 0, rev = 0;
        for (int j = 0; j < w; j++) {
            int c = seq[i + j];
            score += (uint32_t)pwm[c * w + j];
            if (!forward_only) {
                int r = (c == 4) ? 4 : 3 - c;
                rev += (uint32_t)pwm[r * w + (w - j - 1)];
            }
        }
        int32_t max; int k = 0;
        if (forward_only) max = (int32_t)score;
        else {
            uint32_t cdiff = score - rev;
            k = (int)(cdiff >> 31);
            max = (int32_t)(score - (uint32_t)k * cdiff);
        }
        if (max > best) { best = max; pos = (int32_t)i; strand = (int8_t)k; }
    }
    *best_out = best; *pos_out = pos; *strand_out = strand;
}
