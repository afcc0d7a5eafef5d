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
# ignition (reference default tolerances 1e-9 / 1e-15) and its converged twin (1e-11 / 1e-18) through the
# parallel harness ref_bench (same procedure as gen_golden ignite: step loop for the delay, then
# ReactorNet::advance). Sizes per VERDICT r1: >= 512 gri30 / >= 128 nDodecane reactors.
P=${JOBS:-$(nproc)}
rb() { "$B/ref_bench" "$@" > /dev/null; }
ign() { # name yaml phase N seed fuel nO2 O2 N2 Tlo Thi Plo Phi tEnd dT
    local name=$1; shift
    local yaml=$1 phase=$2 n=$3 seed=$4; shift 4
    local dT=${@: -1}; set -- "${@:1:$(($#-1))}"
    rb ignite "$yaml" "$phase" "$n" "$seed" "$P" "$D" "$@" --dump "$G/ignite_$name.ctbm" --tign "$dT"
    rb ignite "$yaml" "$phase" 0 "$seed" "$P" "$D" --states "$G/ignite_$name.ctbm" "$@" --dump "$G/ignite_tight_$name.ctbm" --tign "$dT" --rtol 1e-11 --atol 1e-18
}
ign h2o2 "$D/h2o2.yaml" ohmech 128 $((SEED+10)) H2 0.5 O2 AR 1000 1600 1 20 1e-3 400
ign gri30 "$D/gri30.yaml" gri30 512 $((SEED+12)) CH4 2 O2 N2 1000 2000 1 50 1e-3 400
ign nDodecane "$D/nDodecane_Reitz.yaml" nDodecane_IG 128 $((SEED+13)) c12h26 18.5 o2 n2 700 1400 10 50 1e-2 400
rb ignite "$D/gri30.yaml" gri30 512 $((SEED+14)) "$P" "$D" CH4 2 O2 N2 1200 2200 1 50 1e-5 --dump "$G/substep_gri30.ctbm"
# reactor Jacobian (a13), Jacobian digests at scale, derivative families (f2)
gg rjac "$D/h2o2.yaml" ohmech "$G/rjac_h2o2.ctbm" 32 $((SEED+20))
gg rjac "$D/gri30.yaml" gri30 "$G/rjac_gri30.ctbm" 8 $((SEED+21))
gg rjac "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/rjac_nDodecane.ctbm" 2 $((SEED+22))
gg ddxdigest "$D/h2o2.yaml" ohmech "$G/ddxdigest_h2o2.ctbm" 512 $((SEED+30))
gg ddxdigest "$D/gri30.yaml" gri30 "$G/ddxdigest_gri30.ctbm" 256 $((SEED+31))
gg ddxdigest "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/ddxdigest_nDodecane.ctbm" 64 $((SEED+32))
gg ddxdigest "$TD/pdep-test.yaml" gas "$G/ddxdigest_pdep-test.ctbm" 64 $((SEED+33))
# mixture-averaged transport: fitted polynomials + viscosity / conductivity / mixture diffusion coefficients
gg trans "$D/gri30.yaml" gri30 "$G/trans_gri30.ctbm" 64 20261110
gg trans "$D/h2o2.yaml" ohmech "$G/trans_h2o2.ctbm" 64 20261111
gg derivs "$D/h2o2.yaml" ohmech "$G/derivs_h2o2.ctbm" 32 $((SEED+40))
gg derivs "$D/gri30.yaml" gri30 "$G/derivs_gri30.ctbm" 16 $((SEED+41))
gg derivs "$D/nDodecane_Reitz.yaml" nDodecane_IG "$G/derivs_nDodecane.ctbm" 8 $((SEED+42))
for t in pdep-test:gas sri-falloff:gas tsang-falloff:gas chemically-activated-reaction:gas explicit-third-bodies:gas; do
    gg derivs "$TD/${t%%:*}.yaml" "${t##*:}" "$G/derivs_${t%%:*}.ctbm" 8 $((SEED+43))
done
# other reactor types (f1): RHS and a short ignition history each
for rt in IdealGasMoleReactor IdealGasReactor IdealGasConstPressureReactor; do
    gg rhs "$D/h2o2.yaml" ohmech "$G/rhs_${rt}_h2o2.ctbm" 32 $((SEED+50)) $rt
    gg rhs "$D/gri30.yaml" gri30 "$G/rhs_${rt}_gri30.ctbm" 16 $((SEED+51)) $rt
    rb ignite "$D/h2o2.yaml" ohmech 64 $((SEED+52)) "$P" "$D" H2 0.5 O2 AR 1000 1600 1 20 1e-3 --reactor $rt --dump "$G/ignite_${rt}_h2o2.ctbm" --tign 400
    rb ignite "$D/gri30.yaml" gri30 64 $((SEED+53)) "$P" "$D" CH4 2 O2 N2 1000 2000 1 50 1e-3 --reactor $rt --dump "$G/ignite_${rt}_gri30.ctbm" --tign 400
done
gzip -9 -f -n "$G"/rates_*.ctbm "$G"/ddx_*.ctbm "$G"/rhs_*.ctbm "$G"/ignite_*.ctbm "$G"/substep_*.ctbm \
    "$G"/rjac_*.ctbm "$G"/ddxdigest_*.ctbm "$G"/derivs_*.ctbm "$G"/trans_*.ctbm
ls -la "$G" "$M"
