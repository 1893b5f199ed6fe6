// GERG-2004 / GERG-2008 isochoric family, lane-cooperative form: L lanes of a warp share one state point (32 / L points per warp).
//
// Same mathematics as tq_gerg_iso.cuh (analytic chain rule over the factorised term program; reference: derivs.hpp:1031-1084,
// 1115-1124, 1208-1230, 1264-1437), different work distribution.  The thread-per-point kernel keeps 4 nw + 3 nc staging rows PER THREAD
// (207 rows = 1.6 KB for the 21-component mixture), which caps an SM at 4 warps -- one per scheduler, every dependent FP64 chain and
// every shared-memory round trip exposed (profiles/r02i_ncu_gerg_iso21_summary.json).  Here the staging rows belong to a POINT and the
// L lanes that share it split the work:
//   pass A   rows of the reducing-function gradient: a lane owns components a = rowperm[lane + L m]; for each it walks the row's partner
//            list (general pairs first, beta_T = beta_v = 1 pairs after: a lane-uniform loop structure, rows of similar length run together)
//   pass B0  tables tau^t for the distinct exponents and G_m(delta) for the density groups, entries dealt over the lanes
//   pass B1  per-weight sums a_w, a_w,s, a_w,r: a lane owns weights wperm[lane + L m] (sorted by term count) and walks their term lists
//   pass B2  departure weights -> component rows, by the component's owner through an adjacency list
//   pass C   Hessian: a lane owns rows a = lane + L m and takes the partners b = a + s (mod nc), s = 1 .. (nc-1)/2 -- every unordered pair
//            once, the same number of partners for every row -- and writes H_ab and H_ba
// Sums over components / weights are butterfly reductions over the L lanes.  The per-model tables (pairs in an orientation-free form,
// term lists, groups) are one blob that a CTA copies into shared memory once; the CTA is persistent.
#pragma once
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "tq_common.hpp"
#include "tq_gerg.cuh"

namespace tq {

// pair (a, b) seen from component a: u = ra rb (ra + rb) / (k ra + rb), weight Y.  For a < b: k = beta^2, Y = the reducing pair's
// prefactor; for a > b: k = 1 / beta^2, Y = prefactor / beta^2 (the same function with the roles of the two concentrations swapped)
struct GergIsoPair { double kT, YT, kV, YV; };
struct GergIsoGroup { int kind, d, l, pad_; double c, eta, eps, beta, gamma, bd1, bd2, bl1, bl2; };
struct GergIsoTerm { double n, t; int ti, gi; };
static_assert(sizeof(GergIsoPair) == 32 && sizeof(GergIsoGroup) == 88 && sizeof(GergIsoTerm) == 24, "blob records are whole 8-byte words");

struct GergIsoHdr {   // offsets in 8-byte words from the start of the blob
    int nc, nw, ndep, ngroup, nT, nterm, nwords, pad_;
    int oTc, ovc, opair, opartner, ongen, orowperm, odepF, odepi, odepj, odepq, oadjs, oadjq, oadjo, otval, ogroup, owperm;
    int L, nround, oround, otn, ott, otid;   // term table: lane-interleaved planes for L lanes per point (see gerg_iso_build_blob)
};
static_assert(sizeof(GergIsoHdr) % 8 == 0, "header is whole words");
struct GergIsoRound { int base, len; };   // entries [base + j L + lane], j < len

TQ_HD constexpr int gerg_coop_rows(int nc, int nw, int nT, int ngroup, bool want_h) {
    return want_h ? 4 * nw + 4 * nc + (nc > nT + 3 * ngroup ? nc : nT + 3 * ngroup) : 2 * nw + 2 * nc + nT + 2 * ngroup;
}

// distance between the staging areas of two points of a warp: the row count rounded up to L (mod 16) doubles, so that the groups that share a
// shared-memory wavefront (a half-warp for 8-byte accesses: 16 / L groups) sit on disjoint banks when they run the same row pattern
TQ_HD constexpr int gerg_coop_pstride(int rows, int L) {
    int p = rows;
    while ((p & 15) != (L & 15)) ++p;
    return p;
}

// ---- host: GergProgram -> blob ---------------------------------------------------------------------------------------------------
inline std::vector<double> gerg_iso_build_blob(const GergProgram& P, const int L = 1) {
    const int nc = P.nc, nw = P.nw, ndep = P.ndep;
    struct T3 { int w; double n, t; int g; };
    std::vector<T3> terms;
    std::map<double, int> tix;
    {
        int rec = 0;
        auto widx = [](int byte_off) { return byte_off / (8 * kGergWS); };
        for (int g = 0; g < P.ngroup; ++g)
            for (int q = 0; q < P.group[g].ntq; ++q) {
                const GergRec hd = P.rec[rec];
                const double n0 = P.rec[rec + 1].val;
                rec += 2;
                tix.emplace(hd.val, 0);
                terms.push_back({widx(hd.b), n0, hd.val, g});
                for (int k = 0; k < 2 * hd.a; ++k, ++rec)
                    if (P.rec[rec].val != 0.0) terms.push_back({widx(P.rec[rec].a), P.rec[rec].val, hd.val, g});
            }
    }
    int nT = 0;
    for (auto& kv : tix) kv.second = nT++;
    // weights by descending term count
    std::vector<int> cnt(nw, 0), wperm(nw);
    for (auto& t : terms) cnt[t.w]++;
    for (int w = 0; w < nw; ++w) wperm[w] = w;
    std::stable_sort(wperm.begin(), wperm.end(), [&](int a, int b) { return cnt[a] > cnt[b]; });
    // partner lists and row order
    auto pidx = [&](int i, int j) { return i * nc - i * (i + 1) / 2 + (j - i - 1); };   // i < j, lexicographic
    auto is_b1 = [&](int a, int b) { const GergRedPair& pr = P.pair[a < b ? pidx(a, b) : pidx(b, a)]; return pr.beta2T == 1.0 && pr.beta2V == 1.0; };
    std::vector<int> ngen(nc, 0), rowperm(nc);
    std::vector<unsigned char> partner((size_t)nc * std::max(nc - 1, 1));
    for (int a = 0; a < nc; ++a) {
        int k = 0;
        for (int b = 0; b < nc; ++b) if (b != a && !is_b1(a, b)) partner[(size_t)a * (nc - 1) + k++] = (unsigned char)b;
        ngen[a] = k;
        for (int b = 0; b < nc; ++b) if (b != a && is_b1(a, b)) partner[(size_t)a * (nc - 1) + k++] = (unsigned char)b;
        rowperm[a] = a;
    }
    std::stable_sort(rowperm.begin(), rowperm.end(), [&](int a, int b) { return ngen[a] > ngen[b]; });

    GergIsoHdr h{};
    h.nc = nc; h.nw = nw; h.ndep = ndep; h.ngroup = P.ngroup; h.nT = nT; h.nterm = (int)terms.size();
    int words = sizeof(GergIsoHdr) / 8;
    auto take = [&](size_t bytes) { const int o = words; words += (int)((bytes + 7) / 8); return o; };
    h.oTc = take(8 * nc); h.ovc = take(8 * nc);
    words += words & 1;   // 16-byte alignment of the pair records
    h.opair = take(sizeof(GergIsoPair) * nc * nc);   // four planes kT, YT, kV, YV of nc*nc doubles (structure of arrays: see pair_at)
    h.opartner = take(partner.size()); h.ongen = take(nc); h.orowperm = take(nc);
    h.odepF = take(8 * std::max(ndep, 1)); h.odepi = take(std::max(ndep, 1)); h.odepj = take(std::max(ndep, 1));
    h.odepq = take((size_t)nc * nc);
    h.oadjs = take(nc + 1); h.oadjq = take(std::max(2 * ndep, 1)); h.oadjo = take(std::max(2 * ndep, 1));
    h.otval = take(8 * nT);
    h.ogroup = take(sizeof(GergIsoGroup) * P.ngroup);
    h.owperm = take(nw);
    // Term table.  Lane l of a point walks the weights wperm[l + L m], m = 0, 1, ...: the terms of the L weights of round m are stored interleaved,
    // entry (j, l) at base_m + j L + l, as three planes (n, t, packed indices), the shorter lists padded with n = 0 entries up to the round's longest
    // (the weights are sorted by term count, so little is padded).  The lanes of a group then read CONSECUTIVE shared-memory words in every step
    // (no bank conflicts, no per-lane loop bounds); the 24-byte records of the first version put them on arbitrary banks.
    h.L = L;
    h.nround = (nw + L - 1) / L;
    std::vector<GergIsoRound> rounds((size_t)h.nround);
    int nent = 0;
    for (int m = 0; m < h.nround; ++m) {
        int len = 0;
        for (int l = 0; l < L && m * L + l < nw; ++l) len = std::max(len, cnt[wperm[m * L + l]]);
        rounds[m] = {nent, len};
        nent += len * L;
    }
    h.oround = take(sizeof(GergIsoRound) * h.nround);
    h.otn = take(8 * (size_t)std::max(nent, 1)); h.ott = take(8 * (size_t)std::max(nent, 1)); h.otid = take(4 * (size_t)std::max(nent, 1));
    words += words & 1;
    h.nwords = words;
    std::vector<double> blob((size_t)words, 0.0);
    std::memcpy(blob.data(), &h, sizeof(h));
    auto B = [&](int off) { return reinterpret_cast<unsigned char*>(blob.data() + off); };
    for (int i = 0; i < nc; ++i) { blob[h.oTc + i] = P.Tc[i]; blob[h.ovc + i] = P.vc[i]; }
    // planes instead of records: in pass C the lanes of a group read pairs (a, a + s) for consecutive rows a, i.e. plane indices that differ by
    // nc + 1 -- distinct shared-memory banks for the usual odd / even nc -- where 32-byte records put every lane on one of four bank groups
    for (int a = 0; a < nc; ++a)
        for (int b = 0; b < nc; ++b) {
            GergIsoPair v{1.0, 0.0, 1.0, 0.0};
            if (a != b) {
                const GergRedPair& p = P.pair[a < b ? pidx(a, b) : pidx(b, a)];
                v = (a < b) ? GergIsoPair{p.beta2T, p.YT2, p.beta2V, p.YV2}
                            : GergIsoPair{1.0 / p.beta2T, p.YT2 / p.beta2T, 1.0 / p.beta2V, p.YV2 / p.beta2V};
            }
            const int q = a * nc + b, n2 = nc * nc;
            blob[h.opair + q] = v.kT; blob[h.opair + n2 + q] = v.YT; blob[h.opair + 2 * n2 + q] = v.kV; blob[h.opair + 3 * n2 + q] = v.YV;
        }
    std::memcpy(B(h.opartner), partner.data(), partner.size());
    for (int a = 0; a < nc; ++a) { B(h.ongen)[a] = (unsigned char)ngen[a]; B(h.orowperm)[a] = (unsigned char)rowperm[a]; }
    {
        std::vector<std::vector<std::pair<int, int>>> adj(nc);
        for (int q = 0; q < ndep; ++q) {
            blob[h.odepF + q] = P.depF[q];
            B(h.odepi)[q] = P.dep_i[q]; B(h.odepj)[q] = P.dep_j[q];
            B(h.odepq)[P.dep_i[q] * nc + P.dep_j[q]] = (unsigned char)(q + 1);
            B(h.odepq)[P.dep_j[q] * nc + P.dep_i[q]] = (unsigned char)(q + 1);
            adj[P.dep_i[q]].push_back({q, P.dep_j[q]});
            adj[P.dep_j[q]].push_back({q, P.dep_i[q]});
        }
        int e = 0;
        for (int a = 0; a < nc; ++a) {
            B(h.oadjs)[a] = (unsigned char)e;
            for (auto& qo : adj[a]) { B(h.oadjq)[e] = (unsigned char)qo.first; B(h.oadjo)[e] = (unsigned char)qo.second; ++e; }
        }
        B(h.oadjs)[nc] = (unsigned char)e;
    }
    for (auto& kv : tix) blob[h.otval + kv.second] = kv.first;
    GergIsoGroup* gg = reinterpret_cast<GergIsoGroup*>(blob.data() + h.ogroup);
    for (int g = 0; g < P.ngroup; ++g) {
        const GergGroup& s = P.group[g];
        gg[g] = {s.kind, s.d, s.l, 0, s.c, s.eta, s.eps, s.beta, s.gamma, s.bd[0], s.bd[1], s.bl[0], s.bl[1]};
    }
    std::memcpy(blob.data() + h.oround, rounds.data(), sizeof(GergIsoRound) * rounds.size());
    int* tid = reinterpret_cast<int*>(blob.data() + h.otid);
    for (int p = 0; p < nw; ++p) {
        B(h.owperm)[p] = (unsigned char)wperm[p];
        const int m = p / L, l = p % L;
        int j = 0;
        for (auto& t : terms)
            if (t.w == wperm[p]) {
                const int e = rounds[m].base + j * L + l;
                blob[h.otn + e] = t.n; blob[h.ott + e] = t.t; tid[e] = tix[t.t] | (t.g << 16);
                ++j;
            }
    }
    return blob;
}

// ---- lane-group primitives ---------------------------------------------------------------------------------------------------
template<int L>
TQ_HD double lane_sum(double v) {
#if defined(__CUDA_ARCH__)
#pragma unroll
    for (int s = L / 2; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
#endif
    return v;
}
TQ_HD void lane_sync() {
#if defined(__CUDA_ARCH__)
    __syncwarp();
#endif
}

// One state point, shared by L lanes (lane = 0 .. L-1).  W: the point's staging rows, contiguous (row r at W[r]); rows [0, nc)
// hold the molar concentrations on entry.  valid = false: a padding point (all stores suppressed; the lanes still take part in the
// reductions).  On the host: L = 1.
template<int L, bool WANT_H>
TQ_HD void gerg_iso_coop_point(const double* blob, const double Rgas, const double T, double* W, const int lane, const bool valid,
                               const IsochoricOut& o, const long long k, const long long npts, const IsoPropsOut* ipp = nullptr) {
    const GergIsoHdr& h = *reinterpret_cast<const GergIsoHdr*>(blob);
    const int nc = h.nc, nw = h.nw, ndep = h.ndep;
    auto bytes = [&](int off) { return reinterpret_cast<const unsigned char*>(blob + off); };
#define TQ_ROW(r) W[r]
    const int rA = nw, rS = 2 * nw, rR = 3 * nw;
    const int rvt = WANT_H ? 4 * nw : 2 * nw, rvd = rvt + nc;
    const int rDT = rvd + nc, rDV = rDT + nc;
    const int rX = WANT_H ? rDV + nc : rvd + nc;
    const int rE = rX, rTt = rX, rGt = rX + h.nT;
    constexpr int NG = WANT_H ? 3 : 2;
    const double* Tc = blob + h.oTc;
    const double* vc = blob + h.ovc;
    const double* pairp = blob + h.opair;
    const int n2 = nc * nc;
    auto pair_at = [&](int q) { return GergIsoPair{pairp[q], pairp[n2 + q], pairp[2 * n2 + q], pairp[3 * n2 + q]}; };

    // ---- pass A ------------------------------------------------------------------------------------------------------------------
    double rsum = 0.0, YT1 = 0.0, YV1 = 0.0, YT2 = 0.0, YV2 = 0.0;
    {
        const unsigned char* rowperm = bytes(h.orowperm);
        const unsigned char* ngen = bytes(h.ongen);
        const unsigned char* partner = bytes(h.opartner);
        for (int p = lane; p < nc; p += L) {
            const int a = rowperm[p];
            const double ra = TQ_ROW(a);
            rsum += ra;
            YT1 = fma(ra * ra, Tc[a], YT1);
            YV1 = fma(ra * ra, vc[a], YV1);
            double gT = 2.0 * ra * Tc[a], gV = 2.0 * ra * vc[a], DT = 0.0, DV = 0.0;
            const int pa = a * nc;
            const unsigned char* pl = partner + a * (nc - 1);
            const int ng = ngen[a];
            int idx = 0;
            for (; idx < ng; ++idx) {
                const int b = pl[idx];
                const double rb = TQ_ROW(b);
                const GergIsoPair pr = pair_at(pa + b);
                const double dT = fma(pr.kT, ra, rb), dV = fma(pr.kV, ra, rb);
                const double den = dT * dV;
                if (den == 0.0) continue;   // both concentrations zero: the reference skips the pair (GERG.hpp:189)
                const double inv = fast_rcp(den);
                const double invT = inv * dV, invV = inv * dT;
                const double num = (ra * rb) * (ra + rb), c2 = rb * fma(2.0, ra, rb);
                const double uT = num * invT, uV = num * invV;
                const double uaT = fma(-uT, pr.kT, c2) * invT, uaV = fma(-uV, pr.kV, c2) * invV;
                YT2 = fma(pr.YT, uT, YT2);
                YV2 = fma(pr.YV, uV, YV2);
                gT = fma(pr.YT, uaT, gT);
                gV = fma(pr.YV, uaV, gV);
                if (WANT_H) {
                    DT = fma(pr.YT * invT, 2.0 * fma(-uaT, pr.kT, rb), DT);
                    DV = fma(pr.YV * invV, 2.0 * fma(-uaV, pr.kV, rb), DV);
                }
            }
            for (; idx < nc - 1; ++idx) {   // beta_T = beta_v = 1: u = ra rb (zero, with zero gradient, when the pair is skipped)
                const int b = pl[idx];
                const double rb = TQ_ROW(b);
                const double prYT = pairp[n2 + pa + b], prYV = pairp[3 * n2 + pa + b];
                const double pab = ra * rb;
                YT2 = fma(prYT, pab, YT2);
                YV2 = fma(prYV, pab, YV2);
                gT = fma(prYT, rb, gT);
                gV = fma(prYV, rb, gV);
            }
            TQ_ROW(rvt + a) = gT;
            TQ_ROW(rvd + a) = gV;
            if (WANT_H) { TQ_ROW(rDT + a) = DT; TQ_ROW(rDV + a) = DV; }
        }
    }
    const double rhotot = lane_sum<L>(rsum);
    const double YT = lane_sum<L>(fma(0.5, YT2, YT1)), YV = lane_sum<L>(fma(0.5, YV2, YV1));   // every pair was seen from both ends
    const double invr = 1.0 / rhotot;
    const double iYT = 1.0 / YT, iYV = 1.0 / YV;
    const double tau = YT * (invr * invr) / T;
    const double delta = YV * invr;
    {
        const unsigned char* rowperm = bytes(h.orowperm);
        for (int p = lane; p < nc; p += L) {   // same owner as above: no synchronisation needed
            const int a = rowperm[p];
            TQ_ROW(rvt + a) *= iYT;
            TQ_ROW(rvd + a) *= iYV;
        }
    }
    const double* depF = blob + h.odepF;
    const unsigned char* dep_i = bytes(h.odepi);
    const unsigned char* dep_j = bytes(h.odepj);
    for (int q = lane; q < ndep; q += L) TQ_ROW(nc + q) = TQ_ROW(dep_i[q]) * TQ_ROW(dep_j[q]) * depF[q] * invr;

    // ---- pass B0: tau^t for the distinct exponents, G_m = delta^m/m! d^m/ddelta^m [delta^d exp(h(delta))] for the groups -----------
    double l2tau = log2(tau);
    const double nanfix = l2tau - l2tau;
    l2tau = (l2tau > 30.0) ? 30.0 : l2tau;
    l2tau = (l2tau < -30.0) ? -30.0 : l2tau;
    {
        const Exp2K K = Exp2K::load();
        const double* tval = blob + h.otval;
        for (int u = lane; u < h.nT; u += L) TQ_ROW(rTt + u) = fast_exp2_prod(tval[u], l2tau, K);
        const GergIsoGroup* group = reinterpret_cast<const GergIsoGroup*>(blob + h.ogroup);
        for (int g = lane; g < h.ngroup; g += L) {
            const GergIsoGroup& gr = group[g];
            double E0 = 1.0, E1 = 0.0, E2 = 0.0;
            if (gr.kind != GG_POLY) {
                double h0, h1, h2;
                if (gr.kind == GG_EXP) {
                    const double cdl = -gr.c * ipow(delta, gr.l);
                    h0 = cdl; h1 = gr.bl1 * cdl; h2 = gr.bl2 * cdl;
                } else {
                    const double de = delta - gr.eps;
                    h0 = -gr.eta * (de * de) - gr.beta * (delta - gr.gamma);
                    h1 = delta * (-2.0 * gr.eta * de - gr.beta);
                    h2 = -gr.eta * (delta * delta);
                }
                E0 = fast_exp(h0);
                E1 = h1 * E0;
                E2 = 0.5 * fma(h1, E1, 2.0 * h2 * E0);
            }
            const double dd = ipow(delta, gr.d);
            TQ_ROW(rGt + NG * g) = E0 * dd;
            TQ_ROW(rGt + NG * g + 1) = fma(gr.bd1, E0, E1) * dd;
            if (WANT_H) TQ_ROW(rGt + NG * g + 2) = fma(gr.bd2, E0, fma(gr.bd1, E1, E2)) * dd;
        }
    }
    lane_sync();

    // ---- pass B1: per-weight sums and the totals A[n][m] = Phi_{tau^n delta^m} tau^n delta^m / (n! m!) ----------------------------
    double A00 = 0.0, A10 = 0.0, A01 = 0.0, A20 = 0.0, A11 = 0.0, A02 = 0.0;
    {
        const unsigned char* wperm = bytes(h.owperm);
        const GergIsoRound* round = reinterpret_cast<const GergIsoRound*>(blob + h.oround);
        const double* tn = blob + h.otn;
        const double* tt = blob + h.ott;
        const int* tid = reinterpret_cast<const int*>(blob + h.otid);
        const int nround = h.nround;
        for (int m = 0, p = lane; m < nround; ++m, p += L) {
            double a = 0.0, s = 0.0, r = 0.0, ss = 0.0, sr = 0.0, rr = 0.0;
            const GergIsoRound rd = round[m];
            int e = rd.base + lane;
#pragma unroll 2
            for (int j = 0; j < rd.len; ++j, e += L) {   // the same trip count for every lane: padded entries have n = 0
                const int id = tid[e];
                const double tmt = tt[e];
                const int ti = id & 0xffff, gi = NG * (id >> 16);
                const double x = tn[e] * TQ_ROW(rTt + ti);
                const double y = x * tmt;
                const double G0 = TQ_ROW(rGt + gi), G1 = TQ_ROW(rGt + gi + 1);
                a = fma(x, G0, a);
                s = fma(y, G0, s);
                r = fma(x, G1, r);
                if (WANT_H) {
                    const double G2 = TQ_ROW(rGt + gi + 2);
                    ss = fma(y * fma(0.5, tmt, -0.5), G0, ss);
                    sr = fma(y, G1, sr);
                    rr = fma(x, G2, rr);
                }
            }
            if (p < nw) {
                const int w = wperm[p];
                const double Ww = TQ_ROW(w);
                TQ_ROW(rA + w) = a;
                if (WANT_H) { TQ_ROW(rS + w) = s; TQ_ROW(rR + w) = r; }
                A00 = fma(Ww, a, A00); A10 = fma(Ww, s, A10); A01 = fma(Ww, r, A01);
                if (WANT_H) { A20 = fma(Ww, ss, A20); A11 = fma(Ww, sr, A11); A02 = fma(Ww, rr, A02); }
            }
        }
    }
    A00 = lane_sum<L>(A00) + nanfix; A10 = lane_sum<L>(A10) + nanfix; A01 = lane_sum<L>(A01) + nanfix;
    if (WANT_H) { A20 = lane_sum<L>(A20) + nanfix; A11 = lane_sum<L>(A11) + nanfix; A02 = lane_sum<L>(A02) + nanfix; }
    lane_sync();

    // ---- pass B2: reductions over the departure weights ---------------------------------------------------------------------------
    const double RT = Rgas * T;
    const double Fs = A10, Fr = A01;
    double S0 = 0.0, Ss = 0.0, Sr = 0.0;
    for (int q = lane; q < ndep; q += L) {
        const double wq = TQ_ROW(nc + q), aq = TQ_ROW(rA + nc + q);
        S0 = fma(aq, wq, S0);
        if (WANT_H) { Ss = fma(TQ_ROW(rS + nc + q), wq, Ss); Sr = fma(TQ_ROW(rR + nc + q), wq, Sr); }
        TQ_ROW(rA + nc + q) = depF[q] * aq;   // M_ij
    }
    S0 = lane_sum<L>(S0);
    if (WANT_H) { Ss = lane_sum<L>(Ss); Sr = lane_sum<L>(Sr); }
    lane_sync();
    const double S0r = S0 * invr, Ssr = Ss * invr, Srr = Sr * invr;
    const double invr2 = invr * invr;
    double lnZ = 0.0;
    if (o.lnphi || o.phi) lnZ = log(1.0 + Fr * invr);   // Z = 1 + Ar01, Ar01 = Phi_r / rho (derivs.hpp:1221-1225)
    if (lane == 0 && valid && o.Psir) o.Psir[k] = RT * A00;
    {
        const unsigned char* adjs = bytes(h.oadjs);
        const unsigned char* adjq = bytes(h.oadjq);
        const unsigned char* adjo = bytes(h.oadjo);
        for (int a = lane; a < nc; a += L) {
            double ea = 0.0, es = 0.0, er = 0.0;
            const int e1 = adjs[a + 1];
            for (int e = adjs[a]; e < e1; ++e) {
                const int q = adjq[e];
                const double ro = TQ_ROW(adjo[e]);
                ea = fma(TQ_ROW(rA + nc + q), ro, ea);
                if (WANT_H) { es = fma(depF[q] * TQ_ROW(rS + nc + q), ro, es); er = fma(depF[q] * TQ_ROW(rR + nc + q), ro, er); }
            }
            const double awa = fma(ea, invr, TQ_ROW(rA + a));
            const double tk = TQ_ROW(rvt + a) - 2.0 * invr, dk = TQ_ROW(rvd + a) - invr;
            const double g = fma(Fs, tk, fma(Fr, dk, awa - S0r));
            if (valid) {
                if (o.grad) o.grad[(long long)a * npts + k] = RT * g;
                if (o.lnphi) o.lnphi[(long long)a * npts + k] = g - lnZ;
                if (o.phi) o.phi[(long long)a * npts + k] = exp(g - lnZ);
            }
            if (WANT_H) {
                // all lanes are past their reads of the tau^t table (lane_sync above): E may move into its rows
                TQ_ROW(rA + a) = awa;
                TQ_ROW(rS + a) = fma(es, invr, TQ_ROW(rS + a)) - Ssr;
                TQ_ROW(rR + a) = fma(er, invr, TQ_ROW(rR + a)) - Srr;
                TQ_ROW(rE + a) = -invr2 * ea;
            }
        }
    }
    if (!WANT_H) return;
    lane_sync();

    // ---- pass C: Hessian and the derived properties -------------------------------------------------------------------------------
    const bool want_ip = ipp != nullptr;
    const double Fss = 2.0 * A20 + Fs, Fsr = A11, Frr = 2.0 * A02 + Fr;
    const double Z = 1.0 + Fr * invr;
    const double Ar02 = 2.0 * A02 * invr, Ar11 = Fsr * invr;
    const double dZx0 = Srr + (2.0 * Fsr + Frr) * invr;   // dZ/dx_j = Q_j + dZx0
    if (want_ip && lane == 0 && valid) {
        const IsoPropsOut& ip = *ipp;
        if (ip.pr) ip.pr[k] = RT * Fr;
        if (ip.splus) ip.splus[k] = (A00 - Fs) * invr;
        if (ip.dpdT) ip.dpdT[k] = Rgas * (rhotot + Fr - Fsr);
    }
    const double FsY = Fs * iYT, FrY = Fr * iYV;   // Phi_s / Y_T, Phi_r / Y_V
    const double const0 = (2.0 * S0r + 2.0 * Fs * invr + Fr * invr) * invr;
    const long long ncl = nc;
    const unsigned char* depq = bytes(h.odepq);
    double* const hess = valid ? o.hess : nullptr;
    double* const dlx = (valid && want_ip) ? ipp->dlnphi_dx : nullptr;
    const int nstep = (nc - 1) / 2;
    for (int a = lane; a < nc; a += L) {
        const double ra = TQ_ROW(a);
        const double gTa = TQ_ROW(rvt + a), gVa = TQ_ROW(rvd + a);
        const double ta = gTa - 2.0 * invr, da = gVa - invr;
        const double csa = TQ_ROW(rS + a), cra = TQ_ROW(rR + a), Ea = TQ_ROW(rE + a);
        const double Pa = fma(Fss, ta, fma(Fsr, da, csa)), Qa = fma(Fsr, ta, fma(Frr, da, cra));
        if (want_ip && valid) {
            const IsoPropsOut& ip = *ipp;
            const double ga = fma(Fs, ta, fma(Fr, da, TQ_ROW(rA + a) - S0r));
            const long long oa = (long long)a * npts + k;
            const double R1 = Rgas * (1.0 + log(ra));   // R (rho_ref + ln(rho_i / rho_ref)), rho_ref = 1 (derivs.hpp:1201, 1414)
            if (ip.d2PsirdTdrhoi) ip.d2PsirdTdrhoi[oa] = Rgas * (ga - Pa);
            if (ip.dpdrhovec) ip.dpdrhovec[oa] = fma(RT, Qa, RT);
            if (ip.pmv) ip.pmv[oa] = (1.0 + Qa) / (rhotot * (1.0 + 2.0 * Fr * invr + Ar02));
            if (ip.chempotVLE) ip.chempotVLE[oa] = fma(RT, ga, T * R1);
            if (ip.dchempotdT) ip.dchempotdT[oa] = Rgas * (ga - Pa) + R1;
            if (ip.dlnphi_dT) ip.dlnphi_dT[oa] = (Ar11 / Z - Pa) / T;
            const double dr = (Qa - (Fr * invr + Ar02) / Z) * invr;
            if (ip.dlnphi_drho) ip.dlnphi_drho[oa] = dr;
            if (ip.dlnphi_dv) ip.dlnphi_dv[oa] = -dr * (rhotot * rhotot);
        }
        const double dZxa = Qa + dZx0;
        const int pa = a * nc;
        const int ns = nstep + (((nc & 1) == 0 && a < nc / 2) ? 1 : 0);   // even nc: the antipodal pair belongs to its lower row
        int b = a;
#pragma unroll 2
        for (int s = 1; s <= ns; ++s) {   // two partners in flight: their dependency chains interleave
            b = (b + 1 == nc) ? 0 : b + 1;
            const double rb = TQ_ROW(b);
            const GergIsoPair pr = pair_at(pa + b);
            double yT = 0.0, yV = 0.0;   // Y_ab / Y
            if (pr.kT == 1.0 && pr.kV == 1.0) {
                if (ra + rb != 0.0) { yT = pr.YT * iYT; yV = pr.YV * iYV; }
            } else {
                const double dT = fma(pr.kT, ra, rb), dV = fma(pr.kV, ra, rb);
                const double den = dT * dV;
                if (den != 0.0) {
                    const double inv = fast_rcp(den);
                    const double invT = inv * dV, invV = inv * dT;
                    const double num = (ra * rb) * (ra + rb), c2 = rb * fma(2.0, ra, rb), c3 = ra * fma(2.0, rb, ra), s2 = 2.0 * (ra + rb);
                    const double uT = num * invT, uV = num * invV;
                    const double uaT = fma(-uT, pr.kT, c2) * invT, uaV = fma(-uV, pr.kV, c2) * invV;
                    const double ubT = (c3 - uT) * invT, ubV = (c3 - uV) * invV;
                    yT = pr.YT * iYT * (fma(-ubT, pr.kT, s2 - uaT) * invT);
                    yV = pr.YV * iYV * (fma(-ubV, pr.kV, s2 - uaV) * invV);
                }
            }
            const double gTb = TQ_ROW(rvt + b), gVb = TQ_ROW(rvd + b);
            const double tb = gTb - 2.0 * invr, db = gVb - invr;
            const double crb = TQ_ROW(rR + b);
            double hh = fma(tb, Pa, fma(db, Qa, fma(ta, TQ_ROW(rS + b), da * crb)));
            hh = fma(Fs, fma(-gTa, gTb, yT), hh);
            hh = fma(Fr, fma(-gVa, gVb, yV), hh);
            hh += Ea + TQ_ROW(rE + b) + const0;
            const int q1 = depq[a * nc + b];
            if (q1) hh = fma(TQ_ROW(rA + nc + q1 - 1), invr, hh);
            if (dlx) {
                const double Qb = fma(Fsr, tb, fma(Frr, db, crb));
                dlx[(ncl * a + b) * npts + k] = fma(rhotot, hh, -(Qb + dZx0) / Z);   // d ln phi_a / d x_b
                dlx[(ncl * b + a) * npts + k] = fma(rhotot, hh, -dZxa / Z);          // d ln phi_b / d x_a
            }
            if (hess) {
                hh *= RT;
                hess[(ncl * a + b) * npts + k] = hh;
                hess[(ncl * b + a) * npts + k] = hh;
            }
        }
        {
            const double hda = fma(FsY, TQ_ROW(rDT + a), fma(FrY, TQ_ROW(rDV + a), 2.0 * fma(FsY, Tc[a], FrY * vc[a])));
            double hh = fma(ta, Pa, fma(da, Qa, fma(ta, csa, da * cra)));
            hh = fma(-Fs * gTa, gTa, hh);
            hh = fma(-Fr * gVa, gVa, hh);
            hh += hda + 2.0 * Ea + const0;
            if (dlx) dlx[(ncl * a + a) * npts + k] = fma(rhotot, hh, -dZxa / Z);
            if (hess) hess[(ncl * a + a) * npts + k] = RT * hh;
        }
    }
#undef TQ_ROW
}

// per-model tables on the devices that have asked for them
class GergIsoDeviceTables {
public:
    std::vector<double> host;
    const double* on_device(const GergProgram& P, int dev, int L) const;   // tq_gerg_iso.cu; L: lanes per point the term table is laid out for
    ~GergIsoDeviceTables();
private:
    mutable std::mutex mtx_;
    mutable std::map<int, double*> dev_;
};

// host launchers (tq_gerg_iso.cu); false: the mixture's tables and staging rows do not fit one SM's shared memory (caller falls back)
bool gerg_coop_launch_isochoric(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o,
                                cudaStream_t st);
// both families of outputs from ONE launch (fused C5x step: Psir, gradient, Hessian and the derived properties share passes A, B and C)
bool gerg_coop_launch_fused(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsochoricOut& o, const IsoPropsOut& ip,
                            cudaStream_t st);
bool gerg_coop_launch_isoprops(const GergIsoDeviceTables& tab, const GergProgram& P, double Rgas, const DeviceBatch& in, const IsoPropsOut& o,
                               cudaStream_t st);

}  // namespace tq
