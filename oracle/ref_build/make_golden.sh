#!/bin/bash
# TEST INFRASTRUCTURE — regenerates every committed fixture from the UNMODIFIED reference
# (oracle/_ref, built by build_ref.sh). Container only: needs /root/reference data files.
#   cantera_b200/data/<mech>.ctbm      flat mechanism tables (ctb_extract)
#   tests/golden/<case>.ctbm.gz        seeded inputs + reference outputs (gen_golden)
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
B="$ROOT/oracle/_ref"
D="$B/data"; TD="$B/testdata"
G="$ROOT/tests/golden"; M="$ROOT/cantera_b200/data"
mkdir -p "$G" "$M"
ext() { "$B/ctb_extract" "$1" "$2" "$3" "$D" "$TD"; }
ext "$D/h2o2.yaml" ohmech "$M/h2o2.ctbm"
ext "$D/gri30.yaml" gri30 "$M/gri30.ctbm"
ext "$D/nDodecane_Reitz.yaml" nDodecane_IG "$M/nDodecane_Reitz.ctbm"
for t in pdep-test:gas sri-falloff:gas tsang-falloff:gas jacobian-tests:ohmech \
         chemically-activated-reaction:gas explicit-third-bodies:gas reaction-orders:gas; do
    ext "$TD/${t%%:*}.yaml" "${t##*:}" "$G/mech_${t%%:*}.ctbm"
done
export CANTERA_DATA="$D:$TD"
SEED=20261104
gg() { "$B/gen_golden" "$@"; }
gg rates "$D/h2o2.yaml" ohmech "$G/rates_h2o2.ctbm" 64 $((SEED+0))
gg rates "$D/gri30.yaml" gri30 "$G/rates_gri30.ctbm" 32 $((SEED+1))
gg rates "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/rates_nDodecane.ctbm" 16 $((SEED+3))
for t in pdep-test:gas sri-falloff:gas tsang-falloff:gas jacobian-tests:ohmech \
         chemically-activated-reaction:gas explicit-third-bodies:gas reaction-orders:gas; do
    gg rates "$TD/${t%%:*}.yaml" "${t##*:}" "$G/rates_${t%%:*}.ctbm" 16 $((SEED+7))
    gg ddx "$TD/${t%%:*}.yaml" "${t##*:}" "$G/ddx_${t%%:*}.ctbm" 8 $((SEED+8))
done
gg ddx "$D/h2o2.yaml" ohmech "$G/ddx_h2o2.ctbm" 16 $((SEED+0))
gg ddx "$D/gri30.yaml" gri30 "$G/ddx_gri30.ctbm" 8 $((SEED+1))
gg ddx "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/ddx_nDodecane.ctbm" 3 $((SEED+3))
gg rhs "$D/h2o2.yaml" ohmech "$G/rhs_h2o2.ctbm" 32 $((SEED+0))
gg rhs "$D/gri30.yaml" gri30 "$G/rhs_gri30.ctbm" 32 $((SEED+2))
gg rhs "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/rhs_nDodecane.ctbm" 16 $((SEED+3))
# ignition: fuel nO2 O2 N2 Tlo Thi Plo Phi tEnd dTign   (reference default tolerances 1e-9 / 1e-15)
gg ignite "$D/h2o2.yaml" ohmech "$G/ignite_h2o2.ctbm" 32 $((SEED+0)) H2 0.5 O2 AR 1000 1600 1 20 1e-3 400
gg ignite "$D/gri30.yaml" gri30 "$G/ignite_gri30.ctbm" 48 $((SEED+2)) CH4 2 O2 N2 1000 2000 1 50 1e-3 400
gg ignite "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/ignite_nDodecane.ctbm" 16 $((SEED+3)) c12h26 18.5 o2 n2 700 1400 10 50 1e-2 400
gg ignite "$D/gri30.yaml" gri30 "$G/substep_gri30.ctbm" 32 $((SEED+4)) CH4 2 O2 N2 1200 2200 1 50 1e-5 0
gzip -9 -f -n "$G"/rates_*.ctbm "$G"/ddx_*.ctbm "$G"/rhs_*.ctbm "$G"/ignite_*.ctbm "$G"/substep_*.ctbm
ls -la "$G" "$M"
# converged references (rtol 1e-11 / atol 1e-18): separate the reference's own integration error from ours
gg ignite "$D/h2o2.yaml" ohmech "$G/ignite_tight_h2o2.ctbm" 32 $((SEED+0)) H2 0.5 O2 AR 1000 1600 1 20 1e-3 400 1e-11 1e-18
gg ignite "$D/gri30.yaml" gri30 "$G/ignite_tight_gri30.ctbm" 48 $((SEED+2)) CH4 2 O2 N2 1000 2000 1 50 1e-3 400 1e-11 1e-18
gg ignite "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/ignite_tight_nDodecane.ctbm" 16 $((SEED+3)) c12h26 18.5 o2 n2 700 1400 10 50 1e-2 400 1e-11 1e-18
gzip -9 -f -n "$G"/ignite_tight_*.ctbm
