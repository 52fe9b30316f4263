// model.cpp -- `.hmm` parameters -> log-space state-machine block (host side of SURVEY 8a rows a8/a9).
//
// Replaces hmm_getStateMachine and the four loaders (cPecan/impl/stateMachine.c:476-567, 692-734, 742-764), the
// built-in default machines (:431-468, :663-690) and the emission tables (:218-315).  The state machines are held
// as data -- three ordered transition lists (lower / middle / upper neighbour) -- instead of function pointers;
// the CUDA kernels hard-code the same list structure per state count and take the (eP + tP) sums from DevModel.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>

#include "mab_internal.h"

namespace {

thread_local std::string g_last_error;

enum State { M = 0, SX = 1, SY = 2, LX = 3, LY = 4 };
const double NEG_INF = -std::numeric_limits<double>::infinity();

// the reference's named transition set; a 3-state machine only uses the "short" entries
struct Named {
    double match_continue = 0;
    double from_short_x = 0, from_long_x = 0, open_short_x = 0, extend_short_x = 0, switch_short_x = 0, open_long_x = 0,
           extend_long_x = 0, switch_long_x = 0;
    double from_short_y = 0, from_long_y = 0, open_short_y = 0, extend_short_y = 0, switch_short_y = 0, open_long_y = 0,
           extend_long_y = 0, switch_long_y = 0;
};

void push(mab_model *m, int group, int from, int to, double tp) {
    int k = m->n_trans[group]++;
    m->from[group][k] = from;
    m->to[group][k] = to;
    m->tp[group][k] = tp;
}

void assemble(mab_model *m, const Named &t, const double *em_match16, const double *gap_x4, const double *gap_y4) {
    std::memset(m->n_trans, 0, sizeof(m->n_trans));
    for (int s = 0; s < MAB_MAX_S; s++) m->start[s] = m->ragged_start[s] = m->end[s] = m->ragged_end[s] = 0.0;
    if (m->S == 5) {
        // order of stateMachine5_cellCalculate (stateMachine.c:403-428); its gap-switch transitions are commented out
        push(m, 0, M, SX, t.open_short_x);
        push(m, 0, SX, SX, t.extend_short_x);
        push(m, 0, M, LX, t.open_long_x);
        push(m, 0, LX, LX, t.extend_long_x);
        push(m, 1, M, M, t.match_continue);
        push(m, 1, SX, M, t.from_short_x);
        push(m, 1, SY, M, t.from_short_y);
        push(m, 1, LX, M, t.from_long_x);
        push(m, 1, LY, M, t.from_long_y);
        push(m, 2, M, SY, t.open_short_y);
        push(m, 2, SY, SY, t.extend_short_y);
        push(m, 2, M, LY, t.open_long_y);
        push(m, 2, LY, LY, t.extend_long_y);
        const double end5[5] = {t.match_continue, t.from_short_x, t.from_short_y, t.from_long_x, t.from_long_y};
        const double rend5[5] = {t.open_long_x, t.open_long_x, t.open_long_y, t.extend_long_x, t.extend_long_y};
        for (int s = 0; s < 5; s++) {
            m->start[s] = s == M ? 0.0 : NEG_INF;                             // :350-354
            m->ragged_start[s] = (s == LX || s == LY) ? 0.0 : NEG_INF;        // :356-359
            m->end[s] = end5[s];                                              // :361-378
            m->ragged_end[s] = rend5[s];                                      // :380-397
        }
    } else {
        // order of stateMachine3_cellCalculate (stateMachine.c:642-660)
        push(m, 0, M, SX, t.open_short_x);
        push(m, 0, SX, SX, t.extend_short_x);
        push(m, 0, SY, SX, t.switch_short_x);
        push(m, 1, M, M, t.match_continue);
        push(m, 1, SX, M, t.from_short_x);
        push(m, 1, SY, M, t.from_short_y);
        push(m, 2, M, SY, t.open_short_y);
        push(m, 2, SY, SY, t.extend_short_y);
        push(m, 2, SX, SY, t.switch_short_y);
        const double end3[3] = {t.match_continue, t.from_short_x, t.from_short_y};
        // the match entry is the mean of two LOG values (:627), reproduced as such
        const double rend3[3] = {(t.open_short_x + t.open_short_y) / 2.0, t.extend_short_x, t.extend_short_y};
        for (int s = 0; s < 3; s++) {
            m->start[s] = s == M ? 0.0 : NEG_INF;                             // :595-599
            m->ragged_start[s] = (s == SX || s == SY) ? 0.0 : NEG_INF;        // :601-604
            m->end[s] = end3[s];
            m->ragged_end[s] = rend3[s];
        }
    }
    // N emissions are rounded literals in the reference (:300-315), not log(0.25) / log(1/16)
    for (int x = 0; x < 5; x++) {
        m->em_gap_x[x] = x == 4 ? -1.386294361 : gap_x4[x];
        m->em_gap_y[x] = x == 4 ? -1.386294361 : gap_y4[x];
        for (int y = 0; y < 5; y++)
            m->em_match[x * 5 + y] = (x == 4 || y == 4) ? -2.772588722 : em_match16[x * 4 + y];
    }
}

// emissions_loadGapProbs (:276-298): marginalise the listed gap states' 4x4 blocks onto one axis, normalise, log
void gap_emissions(double out[4], const double *E, const int *x_states, int n_x, const int *y_states, int n_y) {
    for (int i = 0; i < 4; i++) out[i] = 0.0;
    for (int i = 0; i < n_x; i++)
        for (int x = 0; x < 4; x++)
            for (int y = 0; y < 4; y++) out[x] += E[x_states[i] * 16 + x * 4 + y];
    for (int i = 0; i < n_y; i++)
        for (int x = 0; x < 4; x++)
            for (int y = 0; y < 4; y++) out[y] += E[y_states[i] * 16 + x * 4 + y];
    double total = 0.0;
    for (int i = 0; i < 4; i++) total += out[i];
    for (int i = 0; i < 4; i++) out[i] = std::log(out[i] / total);
}

void maybe_swap_short_long(double &ext_s, double &ext_l, double &from_s, double &from_l, double &open_s, double &open_l,
                           double &sw_s, double &sw_l) {
    // :491-497 / :508-514 / :545-551: "long" must be the state with the larger extend probability
    if (ext_s > ext_l) {
        std::swap(ext_s, ext_l);
        std::swap(from_s, from_l);
        std::swap(open_s, open_l);
        std::swap(sw_s, sw_l);
    }
}

}  // namespace

void mab_set_error(const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

extern "C" const char *mab_last_error(void) { return g_last_error.c_str(); }
extern "C" const char *mab_version(void) { return "margin_b200 0.1 (sm_100a, fp64 bit-exact pair-HMM)"; }

extern "C" int mab_model_default(int type, mab_model **out) {
    if (out == nullptr || type < 0 || type > 3) {
        mab_set_error("mab_model_default: bad hmm type %d", type);
        return MAB_E_ARG;
    }
    mab_model *m = new mab_model();
    m->type = type;
    m->S = type <= 1 ? 5 : 3;
    Named t;
    t.match_continue = -0.030064059121770816;
    t.from_short_x = t.from_short_y = -1.272871422049609;
    t.extend_short_x = t.extend_short_y = -0.3388262689231553;
    t.switch_short_x = t.switch_short_y = -4.910694825551255;
    if (m->S == 5) {  // stateMachine5_construct :433-450
        t.from_long_x = t.from_long_y = -5.673280173170473;
        t.open_short_x = t.open_short_y = -4.34381910900448;
        t.open_long_x = t.open_long_y = -6.30810595366929;
        t.extend_long_x = t.extend_long_y = -0.003442492794189331;
        t.switch_long_x = t.switch_long_y = -6.30810595366929;
    } else {  // stateMachine3_construct :665-673
        t.open_short_x = t.open_short_y = -4.21256642;
    }
    const double a = -2.1149196655034745, v = -4.5691014376830479, i = -3.9833860032220842;  // :222-224
    const double em[16] = {a, v, i, v, v, a, v, i, i, v, a, v, v, i, v, a};
    const double g = -1.6094379124341003;  // :238
    const double gap[4] = {g, g, g, g};
    assemble(m, t, em, gap, gap);
    *out = m;
    return MAB_OK;
}

extern "C" int mab_model_from_hmm(int type, const double *T, const double *E, mab_model **out) {
    if (out == nullptr || T == nullptr || E == nullptr || type < 0 || type > 3) {
        mab_set_error("mab_model_from_hmm: bad argument (type %d)", type);
        return MAB_E_ARG;
    }
    mab_model *m = new mab_model();
    m->type = type;
    const int S = m->S = type <= 1 ? 5 : 3;
    const bool symmetric = (type == 0 || type == 2);
    auto t_of = [&](int from, int to) { return T[from * S + to]; };
    auto mean = [&](int f1, int t1, int f2, int t2) { return std::log((t_of(f1, t1) + t_of(f2, t2)) / 2.0); };
    Named t;
    double em[16], gx[4], gy[4];
    for (int x = 0; x < 4; x++)
        for (int y = 0; y < 4; y++) em[x * 4 + y] = std::log(E[M * 16 + x * 4 + y]);  // :247-254
    if (symmetric) {                                                                   // :256-266
        for (int x = 0; x < 4; x++)
            for (int y = x + 1; y < 4; y++)
                em[x * 4 + y] = em[y * 4 + x] = std::log((E[M * 16 + x * 4 + y] + E[M * 16 + y * 4 + x]) / 2.0);
    }
    t.match_continue = std::log(t_of(M, M));
    if (S == 5) {
        const int xs[2] = {SX, LX}, ys[2] = {SY, LY};
        if (!symmetric) {  // stateMachine5_loadAsymmetric
            t.from_short_x = std::log(t_of(SX, M));
            t.from_long_x = std::log(t_of(LX, M));
            t.open_short_x = std::log(t_of(M, SX));
            t.extend_short_x = std::log(t_of(SX, SX));
            t.switch_short_x = std::log(t_of(SY, SX));
            t.open_long_x = std::log(t_of(M, LX));
            t.extend_long_x = std::log(t_of(LX, LX));
            t.switch_long_x = std::log(t_of(LY, LX));
            maybe_swap_short_long(t.extend_short_x, t.extend_long_x, t.from_short_x, t.from_long_x, t.open_short_x,
                                  t.open_long_x, t.switch_short_x, t.switch_long_x);
            t.from_short_y = std::log(t_of(SY, M));
            t.from_long_y = std::log(t_of(LY, M));
            t.open_short_y = std::log(t_of(M, SY));
            t.extend_short_y = std::log(t_of(SY, SY));
            t.switch_short_y = std::log(t_of(SX, SY));
            t.open_long_y = std::log(t_of(M, LY));
            t.extend_long_y = std::log(t_of(LY, LY));
            t.switch_long_y = std::log(t_of(LX, LY));
            maybe_swap_short_long(t.extend_short_y, t.extend_long_y, t.from_short_y, t.from_long_y, t.open_short_y,
                                  t.open_long_y, t.switch_short_y, t.switch_long_y);
            gap_emissions(gx, E, xs, 2, nullptr, 0);
            gap_emissions(gy, E, nullptr, 0, ys, 2);
        } else {  // stateMachine5_loadSymmetric
            t.from_short_x = mean(SX, M, SY, M);
            t.from_long_x = mean(LX, M, LY, M);
            t.open_short_x = mean(M, SX, M, SY);
            t.extend_short_x = mean(SX, SX, SY, SY);
            t.switch_short_x = mean(SX, SY, SY, SX);
            t.open_long_x = mean(M, LX, M, LY);
            t.extend_long_x = mean(LX, LX, LY, LY);
            t.switch_long_x = mean(LX, LY, LY, LX);
            maybe_swap_short_long(t.extend_short_x, t.extend_long_x, t.from_short_x, t.from_long_x, t.open_short_x,
                                  t.open_long_x, t.switch_short_x, t.switch_long_x);
            t.from_short_y = t.from_short_x;
            t.from_long_y = t.from_long_x;
            t.open_short_y = t.open_short_x;
            t.extend_short_y = t.extend_short_x;
            t.switch_short_y = t.switch_short_x;
            t.open_long_y = t.open_long_x;
            t.extend_long_y = t.extend_long_x;
            t.switch_long_y = t.switch_long_x;
            gap_emissions(gx, E, xs, 2, ys, 2);
            gap_emissions(gy, E, xs, 2, ys, 2);
        }
    } else {
        const int xs[1] = {SX}, ys[1] = {SY};
        if (!symmetric) {  // stateMachine3_loadAsymmetric
            t.from_short_x = std::log(t_of(SX, M));
            t.from_short_y = std::log(t_of(SY, M));
            t.open_short_x = std::log(t_of(M, SX));
            t.open_short_y = std::log(t_of(M, SY));
            t.extend_short_x = std::log(t_of(SX, SX));
            t.extend_short_y = std::log(t_of(SY, SY));
            t.switch_short_x = std::log(t_of(SY, SX));
            t.switch_short_y = std::log(t_of(SX, SY));
            gap_emissions(gx, E, xs, 1, nullptr, 0);
            gap_emissions(gy, E, nullptr, 0, ys, 1);
        } else {  // stateMachine3_loadSymmetric
            t.from_short_x = t.from_short_y = mean(SX, M, SY, M);
            t.open_short_x = t.open_short_y = mean(M, SX, M, SY);
            t.extend_short_x = t.extend_short_y = mean(SX, SX, SY, SY);
            t.switch_short_x = t.switch_short_y = mean(SY, SX, SX, SY);
            gap_emissions(gx, E, xs, 1, ys, 1);
            gap_emissions(gy, E, xs, 1, ys, 1);
        }
    }
    assemble(m, t, em, gx, gy);
    *out = m;
    return MAB_OK;
}

extern "C" void mab_model_destroy(mab_model *m) { delete m; }
extern "C" int mab_model_state_number(const mab_model *m) { return m ? m->S : MAB_E_ARG; }
extern "C" int mab_model_type(const mab_model *m) { return m ? m->type : MAB_E_ARG; }

extern "C" int mab_model_transitions(const mab_model *m, double *rows4, int cap_rows) {
    if (m == nullptr) return MAB_E_ARG;
    int n = 0;
    for (int g = 0; g < 3; g++) {
        for (int k = 0; k < m->n_trans[g]; k++, n++) {
            if (rows4 != nullptr && n < cap_rows) {
                rows4[4 * n] = g;
                rows4[4 * n + 1] = m->from[g][k];
                rows4[4 * n + 2] = m->to[g][k];
                rows4[4 * n + 3] = m->tp[g][k];
            }
        }
    }
    return n;
}

extern "C" int mab_model_boundary(const mab_model *m, int which, double *out) {
    if (m == nullptr || out == nullptr || which < 0 || which > 3) return MAB_E_ARG;
    const double *src = which == 0 ? m->start : which == 1 ? m->ragged_start : which == 2 ? m->end : m->ragged_end;
    for (int s = 0; s < m->S; s++) out[s] = src[s];
    return MAB_OK;
}

extern "C" int mab_model_emissions(const mab_model *m, double *em_match25, double *em_gap_x5, double *em_gap_y5) {
    if (m == nullptr) return MAB_E_ARG;
    if (em_match25) std::memcpy(em_match25, m->em_match, sizeof(m->em_match));
    if (em_gap_x5) std::memcpy(em_gap_x5, m->em_gap_x, sizeof(m->em_gap_x));
    if (em_gap_y5) std::memcpy(em_gap_y5, m->em_gap_y, sizeof(m->em_gap_y));
    return MAB_OK;
}

void mab_model_to_device_block(const mab_model *m, DevModel *d) {
    std::memset(d, 0, sizeof(*d));
    d->S = m->S;
    d->n_lower = m->n_trans[0];
    d->n_middle = m->n_trans[1];
    d->n_upper = m->n_trans[2];
    for (int c = 0; c < 5; c++) {
        for (int k = 0; k < m->n_trans[0]; k++) d->etp_lower[c][k] = m->em_gap_x[c] + m->tp[0][k];
        for (int k = 0; k < m->n_trans[2]; k++) d->etp_upper[c][k] = m->em_gap_y[c] + m->tp[2][k];
    }
    for (int c = 0; c < 25; c++)
        for (int k = 0; k < m->n_trans[1]; k++) d->etp_middle[c][k] = m->em_match[c] + m->tp[1][k];
    for (int s = 0; s < MAB_MAX_S; s++) {
        d->start[s] = m->start[s];
        d->ragged_start[s] = m->ragged_start[s];
        d->end[s] = m->end[s];
        d->ragged_end[s] = m->ragged_end[s];
    }
    for (int g = 0; g < 3; g++)
        for (int k = 0; k < m->n_trans[g]; k++) d->tp[g][k] = m->tp[g][k];
    for (int c = 0; c < 25; c++) d->em_match[c] = m->em_match[c];
    for (int c = 0; c < 5; c++) {
        d->em_gap_x[c] = m->em_gap_x[c];
        d->em_gap_y[c] = m->em_gap_y[c];
    }
    int n = 0;
    for (int g = 0; g < 3; g++)
        for (int k = 0; k < m->n_trans[g]; k++, n++) {
            d->t_from[n] = m->from[g][k];
            d->t_to[n] = m->to[g][k];
        }
}
