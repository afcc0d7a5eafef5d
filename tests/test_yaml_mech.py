"""cantera_b200/yaml_mech.py — the stand-alone YAML loader on the CPU:
(1) a hand-written toy input (tests/golden/toy_mech.yaml) against values worked out by hand;
(2) every mechanism the repository holds reference-extracted tables for (cantera_b200/data/*.ctbm,
    tests/golden/mech_*.ctbm: written by oracle/_ref/ctb_extract = the reference's own newSolution()):
    the loader's tables must be IDENTICAL, bit for bit, strings included. The YAML inputs are the
    reference's data files under oracle/_ref (or /root/reference); without them (2) is skipped."""
import math
import os

import numpy as np
import pytest

from conftest import GOLDEN, MECHS, ROOT

from cantera_b200.ctbm import read_arrays
from cantera_b200.yaml_mech import (GAS_CONSTANT, MechanismError, UnitSystem, _detect_third_body, _parse_equation,
                                    tables_from_yaml, unit_factor)

TOY = os.path.join(GOLDEN, "toy_mech.yaml")


def test_unit_factors():
    assert unit_factor("cm") == 0.01 and unit_factor("kmol") == 1.0 and unit_factor("mol") == 1e-3
    assert unit_factor("cal/mol") == 4.184 * math.pow(1e-3, -1.0)
    assert unit_factor("kJ/mol") == 1e3 * math.pow(1e-3, -1.0)
    assert unit_factor("cm^3/mol/s") == math.pow(0.01, 3.0) * math.pow(1e-3, -1.0) * 1.0
    assert unit_factor("atm") == 101325.0 and unit_factor("MPa") == 1e6
    with pytest.raises(MechanismError):
        unit_factor("furlong")
    u = UnitSystem({"length": "cm", "quantity": "mol", "activation-energy": "cal/mol"})
    assert u.activation_energy_over_R(1000.0) == 1000.0 * 4184.0 / GAS_CONSTANT
    assert u.activation_energy_over_R("300 K") == 300.0 * GAS_CONSTANT / GAS_CONSTANT
    assert u.pressure_pa("0.5 bar") == 0.5e5 and u.pressure_pa(3.0) == 3.0
    # no explicit activation energy: energy / quantity of the system
    assert UnitSystem({"quantity": "mol"}).activation_energy_over_R(8.0) == 8.0 * (1.0 / 1e-3) / GAS_CONSTANT


@pytest.mark.parametrize("eq,kind,third,reac,prod,rev", [
    ("2 O + M <=> O2 + M", "Arrhenius", "M", {"O": 2.0}, {"O2": 1.0}, True),
    ("H + O2 (+M) <=> HO2 (+M)", "falloff", "(+M)", {"H": 1.0, "O2": 1.0}, {"HO2": 1.0}, True),
    ("H + O2 (+ M) <=> HO2 (+ M)", "falloff", "(+M)", {"H": 1.0, "O2": 1.0}, {"HO2": 1.0}, True),
    ("H + O2 + H2O <=> HO2 + H2O", "Arrhenius", "H2O", {"H": 1.0, "O2": 1.0}, {"HO2": 1.0}, True),
    ("H + 2 O2 <=> HO2 + O2", "Arrhenius", "O2", {"H": 1.0, "O2": 1.0}, {"HO2": 1.0}, True),
    ("H + O2 + H2O <=> HO2 + H2O", "elementary", None, {"H": 1.0, "O2": 1.0, "H2O": 1.0}, {"HO2": 1.0, "H2O": 1.0}, True),
    ("H2 + O => H + OH", "Arrhenius", None, {"H2": 1.0, "O": 1.0}, {"H": 1.0, "OH": 1.0}, False),
    ("A + B <=> C + A", "Arrhenius", None, {"A": 1.0, "B": 1.0}, {"A": 1.0, "C": 1.0}, True),  # no side has three species
    ("0.5 A + B + C = 0.5 A + D", "Arrhenius", None, {"A": 0.5, "B": 1.0, "C": 1.0}, {"A": 0.5, "D": 1.0}, True),
    ("R1A + R1B (+ AR) <=> P1 (+ AR)", "falloff", "(+AR)", {"R1A": 1.0, "R1B": 1.0}, {"P1": 1.0}, True),
])
def test_equation_and_third_body_detection(eq, kind, third, reac, prod, rev):
    """The cases of Reaction::setEquation (Reaction.cpp:389-555)."""
    r, p, reversible = _parse_equation(eq)
    assert _detect_third_body(r, p, kind, {}, eq) == third
    assert (r, p, reversible) == (reac, prod, rev)


def test_toy_mechanism_by_hand():
    t = tables_from_yaml(TOY, "toy")
    assert list(t["dims"]) == [5, 6]
    assert t["species_names"].split("\n") == ["A2", "A", "B", "NO", "AB"]     # NO stays a string
    # Phase::addSpecies sums in phase element order (O, H, N)
    assert list(t["mol_weights"]) == [2 * 1.008, 1.008, 15.999, 15.999 + 14.007, 15.999 + 1.008]
    assert list(t["nasa_tmid"]) == [1000.0, 1000.0, 1000.0, 6000.0, 1000.0]   # single range: Tmid = Tmax, data repeated
    assert np.array_equal(t["nasa_low"][21:28], t["nasa_high"][21:28])
    assert list(t["rxn_type"]) == [0, 0, 2, 0, 5, 6]
    assert list(t["rxn_flags"]) == [7, 7, 3, 0, 1, 1]
    def si(A, conc):  # (cm, mol) -> (m, kmol) in the multiplication order of UnitSystem::convertTo
        return A / 1.0 * 1.0 * math.pow(0.01, -3.0 * conc) * 1.0 * math.pow(1e-3, conc) * 1.0 * 1.0
    assert t["rate_A"][0] == 1.0e18 / 1.0 * 1.0 * math.pow(0.01, 6.0) * 1.0 * math.pow(1e-3, -2.0) * 1.0 * 1.0
    assert t["rate_A"][1] == si(5e17, -2.0) and t["rate_Ea_R"][1] == 1000.0 * 4184.0 / GAS_CONSTANT
    assert t["rate_A"][2] == si(7.4e13, -1.0) and t["low_A"][2] == si(2.3e18, -2.0)
    assert list(t["falloff_coeffs"][10:15]) == [0.7346, 1.0 / 94.0, 1.0 / 1756.0, 5182.0, 0.0]
    conc = (1.0 - 0.8) - 1.0                                                  # the reference's accumulation order
    assert t["rate_A"][3] == 3.87e4 / 1.0 * 1.0 * math.pow(0.01, -3.0 * conc) * 1.0 * math.pow(1e-3, conc) * 1.0 * 1.0
    assert t["rate_Ea_R"][3] == 26.19 * (1e3 * math.pow(1e-3, -1.0)) / (GAS_CONSTANT * 1.0)
    # generic collider with efficiencies (the undeclared XX is dropped), specific collider, falloff collider
    assert list(t["tb_default"]) == [1.0, 0.0, 1.0, 0.0, 0.0, 0.0]
    assert list(t["tb_ptr"]) == [0, 2, 3, 4, 4, 4, 4]
    assert list(t["tb_species"]) == [0, 3, 3, 0] and list(t["tb_eff"]) == [2.5, 0.5, 1.0, 2.0]
    # stoichiometry in name order; orders beside it
    assert list(t["reac_ptr"]) == [0, 1, 3, 5, 7, 9, 11]
    assert list(t["reac_species"]) == [1, 1, 2, 1, 2, 0, 2, 1, 4, 1, 2]
    assert list(t["reac_stoich"][:1]) == [2.0] and list(t["reac_order"][5:7]) == [0.8, 1.0]
    assert list(t["delta_n"]) == [-1.0, -1.0, -1.0, 0.0, 0.0, -1.0]
    # Plog: sorted by pressure, duplicates kept in input order, P as exp(log P)
    assert list(t["plog_ptr"]) == [0, 0, 0, 0, 0, 3, 3]
    assert list(t["plog_P"]) == [math.exp(math.log(0.01 * 101325.0)), math.exp(math.log(101325.0))] * 1 + [math.exp(math.log(101325.0))]
    assert list(t["plog_A"]) == [si(1.2e14, -1.0), si(4.9e15, -1.0), si(1.0e11, -1.0)]
    # Chebyshev with a local `units: {quantity: molec}`: log10 of the unit offset added to the first coefficient
    offset = 1.0 / 1.0 * 1.0 * math.pow(0.01, 3.0) * 1.0 * math.pow(1.0 / 6.02214076e26, -1.0) * 1.0 * 1.0
    assert list(t["cheb_coeffs"]) == [8.0 + math.log10(offset), 0.5, -1.0, 0.1]
    assert list(t["cheb_limits"][20:24]) == [290.0, 3000.0, 0.01 * 101325.0, 100.0 * 101325.0]
    assert t["equations"].split("\n")[:3] == ["2 A + M <=> A2 + M", "A + B + NO <=> AB + NO", "A + B (+M) <=> AB (+M)"]


def test_declared_species_rule_and_errors():
    t = tables_from_yaml(TOY, "toy-declared")   # only the reaction whose species are all declared survives
    assert list(t["dims"]) == [3, 1] and t["equations"] == "2 A + M <=> A2 + M"
    assert list(t["tb_species"]) == [0, 2] and list(t["tb_eff"]) == [2.5, 0.5]
    with pytest.raises(MechanismError, match="phase 'nope' not found"):
        tables_from_yaml(TOY, "nope")


def _inputs():
    for base in (os.path.join(ROOT, "oracle", "_ref"), None):
        d, td = (os.path.join(base, "data"), os.path.join(base, "testdata")) if base else \
            ("/root/reference/data", "/root/reference/test/data")
        if os.path.exists(os.path.join(d, "gri30.yaml")):
            return d, td
    return None, None


CASES = [("h2o2", "h2o2.yaml", "ohmech"), ("gri30", "gri30.yaml", "gri30"), ("nDodecane", "nDodecane_Reitz.yaml", "nDodecane_IG"),
         ("pdep-test", None, "gas"), ("sri-falloff", None, "gas"), ("tsang-falloff", None, "gas"), ("jacobian-tests", None, "ohmech"),
         ("chemically-activated-reaction", None, "gas"), ("explicit-third-bodies", None, "gas"), ("reaction-orders", None, "gas")]


@pytest.mark.parametrize("case,fname,phase", CASES)
def test_tables_identical_to_the_reference_extractor(case, fname, phase):
    d, td = _inputs()
    if d is None:
        pytest.skip("the reference's YAML data files are not on this machine (oracle/_ref/data)")
    path = os.path.join(d, fname) if fname else os.path.join(td, case + ".yaml")
    ours = tables_from_yaml(path, phase, data_dirs=[d, td])
    ref = read_arrays(MECHS[case])
    assert set(ours) == set(ref)
    for k, v in ref.items():
        if isinstance(v, str):
            assert ours[k] == v, k
        else:
            assert ours[k].dtype == v.dtype and ours[k].shape == v.shape, k
            assert np.array_equal(ours[k], v), "%s: %d of %d entries differ" % (k, int((ours[k] != v).sum()), v.size)


_HEAD = """
units: {length: cm, quantity: mol, activation-energy: cal/mol}
phases:
- name: p
  thermo: ideal-gas
  elements: [H, O]
  species: [A, B, C]
  kinetics: gas
  reactions: all
%s
species:
- {name: A, composition: {H: 1}, thermo: {model: NASA7, temperature-ranges: [200.0, 3000.0], data: [[2.5, 0, 0, 0, 0, 0, 0]]}}
- {name: B, composition: {O: 1}, thermo: {model: NASA7, temperature-ranges: [200.0, 3000.0], data: [[2.5, 0, 0, 0, 0, 0, 0]]}}
- {name: C, composition: {H: 1, O: 1}, thermo: {model: NASA7, temperature-ranges: [200.0, 3000.0], data: [[3.5, 0, 0, 0, 0, 0, 0]]}}
reactions:
%s
"""


def _doc(tmp_path, reactions, phase_extra=""):
    p = tmp_path / "m.yaml"
    p.write_text(_HEAD % (phase_extra, reactions))
    return str(p)


@pytest.mark.parametrize("reactions,match", [
    ("- {equation: A + B <=> C, rate-constant: [1e10, 0, 0]}\n- {equation: A + B <=> C, rate-constant: [2e10, 0, 0]}", "undeclared duplicate"),
    ("- {equation: A + B <=> C, rate-constant: [1e10, 0, 0]}\n- {equation: C <=> A + B, rate-constant: [2e10, 0, 0]}", "undeclared duplicate"),
    ("- {equation: A + B <=> C, rate-constant: [1e10, 0, 0], duplicate: true}", "no duplicate found"),
    ("- {equation: A + B <=> C, rate-constant: [-1e10, 0, 0]}", "negative pre-exponential"),
    ("- {equation: A + B <=> C, rate-constant: [1e10 cm^6/mol^2/s, 0, 0]}", "incompatible units"),
    ("- {equation: A + B <=> C, rate-constant: [1e10, 0, 5 m]}", "activation energy"),
    ("- {equation: A + B => C, rate-constant: [1e10, 0, 0], orders: {C: 1.0}}", "non-reactant"),
    ("- {equation: A + B <=> C, rate-constant: [1e10, 0, 0], efficiencies: {A: 2.0}}", "no third body"),
    ("- {equation: A + B + M <=> C + M, rate-constant: [1e10, 0, 0], efficiencies: {ZZ: 2.0}}", "undeclared species"),
    ("- equation: A + B <=> C\n  type: pressure-dependent-Arrhenius\n  rate-constants:\n  - {P: 1 atm, A: -1e10, b: 0, Ea: 0}\n  - {P: 10 atm, A: 1e10, b: 0, Ea: 0}", "invalid rate coefficient"),
    ("- {equation: A + B <=> C, type: Blowers-Masel, rate-constant: [1e10, 0, 0, 0]}", "not on the batched path"),
])
def test_inputs_the_reference_rejects(tmp_path, reactions, match):
    """Reaction::check, ArrheniusBase::check, UnitSystem::convert, PlogRate::validate, Kinetics::checkDuplicates."""
    with pytest.raises(MechanismError, match=match):
        tables_from_yaml(_doc(tmp_path, reactions), "p")


def test_duplicates_and_explicit_third_body_rule(tmp_path):
    ok = "- {equation: A + B <=> C, rate-constant: [1e10, 0, 0], duplicate: true}\n- {equation: A + B <=> C, rate-constant: [2e10, 0, 1000], duplicate: true}"
    assert list(tables_from_yaml(_doc(tmp_path, ok), "p")["dims"]) == [3, 2]
    # irreversible reactions in opposite directions are no duplicates
    opp = "- {equation: A + B => C, rate-constant: [1e10, 0, 0]}\n- {equation: C => A + B, rate-constant: [2e10, 0, 0]}"
    assert list(tables_from_yaml(_doc(tmp_path, opp), "p")["dims"]) == [3, 2]
    # a generic and a specific collider with a common third body: untouched by default, efficiency zeroed on request
    tb = "- {equation: A + B + M <=> C + M, type: three-body, rate-constant: [1e16, 0, 0], efficiencies: {C: 3.0}}\n" \
         "- {equation: A + B + C <=> C + C, rate-constant: [2e16, 0, 0]}"
    t = tables_from_yaml(_doc(tmp_path, tb), "p")
    assert list(t["tb_species"]) == [2, 2] and list(t["tb_eff"]) == [3.0, 1.0]
    t = tables_from_yaml(_doc(tmp_path, tb, "  explicit-third-body-duplicates: modify-efficiency"), "p")
    assert list(t["tb_species"]) == [2, 2] and list(t["tb_eff"]) == [0.0, 1.0]


def test_every_ideal_gas_phase_of_the_reference_inputs(tmp_path):
    """All YAML files of the reference's data and test-data trees: wherever the reference's own loader + extractor
    (oracle/_ref/ctb_extract) accept an ideal-gas phase, the loader's tables are identical; wherever they refuse
    it, the loader refuses it too."""
    import glob
    import subprocess
    from cantera_b200.yaml_mech import _load
    ext = os.path.join(ROOT, "oracle", "_ref", "ctb_extract")
    D, TD = "/root/reference/data", "/root/reference/test/data"
    if not (os.path.exists(ext) and os.path.isdir(D)):
        pytest.skip("needs the reference tree and oracle/_ref/ctb_extract")
    same = refused = 0
    for path in sorted(glob.glob(D + "/*.yaml") + glob.glob(TD + "/*.yaml")):
        try:
            root = _load(path)
        except Exception:  # noqa: BLE001  (inputs that are not valid YAML on purpose)
            continue
        for ph in (root.get("phases") or []) if isinstance(root, dict) else []:
            if ph.get("thermo") != "ideal-gas" or ph.get("kinetics") not in ("gas", "bulk"):
                continue
            name, out = str(ph["name"]), str(tmp_path / "t.ctbm")
            r = subprocess.run([ext, path, name, out, D, TD], capture_output=True, text=True, timeout=300)
            where = "%s:%s" % (os.path.basename(path), name)
            if r.returncode != 0:
                with pytest.raises(Exception):
                    tables_from_yaml(path, name, data_dirs=[D, TD])
                refused += 1
                continue
            ref, ours = read_arrays(out), tables_from_yaml(path, name, data_dirs=[D, TD])
            for k, v in ref.items():
                if isinstance(v, str):
                    assert ours[k] == v, (where, k)
                else:
                    assert ours[k].shape == v.shape and np.array_equal(ours[k], v, equal_nan=True), (where, k)
            same += 1
    assert same >= 60 and refused >= 20
