// genericSAFT with a PC-SAFT non-polar term [+ association] (/root/reference/include/teqp/models/saft/genericsaft.hpp:20-28, 65-73:
// alphar = nonpolar.alphar + association.alphar): the PC-SAFT model of tq_pcsaft.cuh and the association term of tq_assoc.cuh over the same jets.
#pragma once
#include "tq_assoc.cuh"
#include "tq_pcsaft.cuh"

namespace tq {

struct PcAssocParams {
    PcsaftParams pc;
    AssocParams assoc;
    int has_assoc, pad_;
};

struct PcAssocDev {
    using Params = PcAssocParams;
    static constexpr int kMaxC = kAssocMaxC;   // the association tables bound the component count (<= kPcsaftMaxC)
    static constexpr int kCrossOrder = 1;
    static constexpr bool kNoSmallVariants = true;
    template<int MAXN, class TT, class RT, class XT>
    static TQ_HD prom_t<prom_t<TT, RT>, XT> alphar(const Params& p, const TT& T, const RT& rho, const XT* x) {
        auto a = pcsaft_alphar<MAXN>(p.pc, T, rho, x);
        if (p.has_assoc) a = a + assoc_alphar<MAXN>(p.assoc, p.pc.nc, T, rho, x);
        return a;
    }
    template<class XT> static TQ_HD double R(const Params& p, const XT*) { return p.pc.R; }
};
static_assert(kAssocMaxC <= kPcsaftMaxC, "PC-SAFT arrays must hold every component the association term can");

}  // namespace tq
