"""Stand-alone YAML -> flat mechanism tables: the reference's input format read
WITHOUT the reference, for hosts where libcantera is not available at run time.

    tables = tables_from_yaml("gri30.yaml", "gri30")          # {name: ndarray | str}, the .ctbm keys
    kin = BatchedGasKinetics.from_yaml_standalone("gri30.yaml", "gri30")

The result is what ``tablesFromSolution`` (include/cantera_b200/MechanismExtract.h) produces from a
Solution loaded by the reference's own ``newSolution()``: ``tests/test_yaml_mech.py`` compares every
table with the reference-extracted ``.ctbm`` files (integers and structure exactly, doubles to a few
ulp) for gri30, h2o2, nDodecane_Reitz and the reference's kinetics test mechanisms. The semantics
restated here (each function cites the lines it follows):

* unit systems and conversion arithmetic      src/base/Units.cpp:134-200,562-571,618-632,713-726
* reaction equations, third-body detection    src/kinetics/Reaction.cpp:284-320,389-555,777-830,889-956
* rate-coefficient units                      src/kinetics/Reaction.cpp:576-617, Falloff.cpp:130-161
* Arrhenius / falloff / Plog / Chebyshev       Arrhenius.cpp:35-74, Falloff.cpp:235-270,302-322,405-428,
                                              496-512, PlogRate.cpp:62-74, ChebyshevRate.cpp:63-94
* NASA7 species thermo, molecular weights     SpeciesThermoFactory.cpp:88-106, Phase.cpp:786-845
* species / reaction selection of a phase     docs of the YAML format (phases: species, reactions)

Scope: the ideal-gas phase model with bulk kinetics and the rate types of include/ctb_clib.h
(CTB_RATE_*); anything else raises MechanismError, as the extractor does.
"""
import math
import os
import re

import numpy as np
import yaml

GAS_CONSTANT = 6.02214076e26 * 1.380649e-23  # ct_defs.h:84-120 (a computed product, not a literal)
AVOGADRO = 6.02214076e26
ONE_ATM = 101325.0
SMALL_NUMBER = 1e-300

RATE_ARRHENIUS, RATE_LINDEMANN, RATE_TROE, RATE_SRI, RATE_TSANG, RATE_PLOG, RATE_CHEBYSHEV = range(7)
RXN_REVERSIBLE, RXN_THIRD_BODY, RXN_MASS_ACTION, RXN_CHEM_ACTIVATED = 1, 2, 4, 8

#: standard atomic weights of src/thermo/Elements.cpp:60-200 (IUPAC values; the subset gas mechanisms use)
ATOMIC_WEIGHTS = {
    "H": 1.008, "D": 2.0141017781, "Tr": 3.0160492820, "He": 4.002602, "Li": 6.94, "B": 10.81, "C": 12.011,
    "N": 14.007, "O": 15.999, "F": 18.998403163, "Ne": 20.1797, "Na": 22.98976928, "Mg": 24.305,
    "Al": 26.9815384, "Si": 28.085, "P": 30.973761998, "S": 32.06, "Cl": 35.45, "Ar": 39.95, "K": 39.0983,
    "Ca": 40.078, "Ti": 47.867, "Fe": 55.845, "Ni": 58.6934, "Cu": 63.546, "Zn": 65.38, "Br": 79.904,
    "Kr": 83.798, "E": 9.1093837015e-31 * AVOGADRO,
}


class MechanismError(ValueError):
    pass


# ------------------------------------------------------------------------------------------ YAML

class _Loader(yaml.SafeLoader):
    """YAML 1.2 core-schema booleans only: the species NO, the element Y, ... stay strings (the reference
    reads with yaml-cpp and converts on demand)."""


_Loader.yaml_implicit_resolvers = {k: [(tag, rx) for tag, rx in v if tag != "tag:yaml.org,2002:bool"]
                                   for k, v in yaml.SafeLoader.yaml_implicit_resolvers.items()}
_Loader.add_implicit_resolver("tag:yaml.org,2002:bool", re.compile(r"^(?:true|True|TRUE|false|False|FALSE)$"), list("tTfF"))


def _load(path):
    with open(path) as f:
        return yaml.load(f, Loader=_Loader)


_NUM = re.compile(r"^\s*([+-]?(?:\d+\.?\d*|\.\d+)(?:[eEdD][+-]?\d+)?)\s*(.*?)\s*$")


def _split_unit(v):
    """AnyValue -> (number, unit string) (Units.cpp split_unit): "1.0 atm", 3.2, "5e+11"."""
    if isinstance(v, bool):
        raise MechanismError("expected a number, got %r" % v)
    if isinstance(v, (int, float)):
        return float(v), ""
    m = _NUM.match(str(v))
    if not m:
        raise MechanismError("expected a number (with optional units), got %r" % (v,))
    return float(m.group(1).replace("d", "e").replace("D", "e")), m.group(2)


def _num(v):
    x, u = _split_unit(v)
    if u:
        raise MechanismError("unexpected units in %r" % (v,))
    return x


# ----------------------------------------------------------------------------------------- units

# factor, and the class of the unit when used as an activation energy ("E": energy, "K", "eV")
_KNOWN = {
    "": 1.0, "1": 1.0, "m": 1.0, "g": 1e-3, "s": 1.0, "min": 60.0, "hr": 3600.0, "K": 1.0, "kmol": 1.0, "mol": 1e-3,
    "gmol": 1e-3, "molec": 1.0 / AVOGADRO, "J": 1.0, "cal": 4.184, "erg": 1e-7, "eV": 1.602176634e-19,
    "N": 1.0, "dyn": 1e-5, "W": 1.0, "Pa": 1.0, "atm": ONE_ATM, "bar": 1.0e5, "dyn/cm^2": 0.1,
    "micron": 1e-6, "angstrom": 1e-10, "Å": 1e-10, "L": 1e-3, "l": 1e-3,
}
_PREFIX = {"Y": 1e24, "Z": 1e21, "E": 1e18, "P": 1e15, "T": 1e12, "G": 1e9, "M": 1e6, "k": 1e3, "h": 1e2, "d": 1e-1,
           "c": 1e-2, "m": 1e-3, "u": 1e-6, "n": 1e-9, "p": 1e-12, "f": 1e-15, "a": 1e-18, "z": 1e-21, "y": 1e-24}


# dimensions (length, mass, time, quantity, temperature) of the known units, for the convertibility checks of
# UnitSystem::convert / convertActivationEnergy (Units.cpp:202-211,590-632,676-706)
_DIMS = {
    "": (0, 0, 0, 0, 0), "1": (0, 0, 0, 0, 0), "m": (1, 0, 0, 0, 0), "g": (0, 1, 0, 0, 0), "s": (0, 0, 1, 0, 0),
    "min": (0, 0, 1, 0, 0), "hr": (0, 0, 1, 0, 0), "K": (0, 0, 0, 0, 1), "kmol": (0, 0, 0, 1, 0), "mol": (0, 0, 0, 1, 0),
    "gmol": (0, 0, 0, 1, 0), "molec": (0, 0, 0, 1, 0), "J": (2, 1, -2, 0, 0), "cal": (2, 1, -2, 0, 0), "erg": (2, 1, -2, 0, 0),
    "eV": (2, 1, -2, 0, 0), "N": (1, 1, -2, 0, 0), "dyn": (1, 1, -2, 0, 0), "W": (2, 1, -3, 0, 0), "Pa": (-1, 1, -2, 0, 0),
    "atm": (-1, 1, -2, 0, 0), "bar": (-1, 1, -2, 0, 0), "dyn/cm^2": (-1, 1, -2, 0, 0), "micron": (1, 0, 0, 0, 0),
    "angstrom": (1, 0, 0, 0, 0), "Å": (1, 0, 0, 0, 0), "L": (3, 0, 0, 0, 0), "l": (3, 0, 0, 0, 0),
}


def unit_dims(name):
    """(length, mass, time, quantity, temperature) exponents of a unit string."""
    dims = [0.0] * 5
    body = re.sub(r"^[+-]?(?:[0-9]+(?:[.][0-9]*)?|[.][0-9]+)", "", name)
    pos = 0
    for m in re.finditer(r"[*/]|$", body):
        group = body[pos:m.start()]
        sign = -1.0 if pos and body[pos - 1] == "/" else 1.0
        pos = m.end()
        exponent = 1.0
        if "^" in group:
            group, e = group.split("^", 1)
            exponent = float(e)
        unit = group.strip()
        d = _DIMS[unit] if unit in _DIMS else (_DIMS[unit[1:]] if unit[:1] in _PREFIX and unit[1:] in _DIMS else None)
        if d is None:
            raise MechanismError("unknown unit '%s' in unit string '%s'" % (unit, name))
        for i in range(5):
            dims[i] += sign * exponent * d[i]
        if m.start() == len(body):
            break
    return tuple(dims)


def _same_dims(a, b):
    return all(abs(x - y) < 1e-9 for x, y in zip(a, b))


def unit_factor(name):
    """Conversion factor of a unit string to SI + kmol, with the arithmetic of Units::Units(string)
    (Units.cpp:134-200): group by group, factor *= pow(prefix * unit, +-exponent)."""
    factor = 1.0
    m = re.match(r"[+-]?(?:[0-9]+(?:[.][0-9]*)?|[.][0-9]+)", name)
    start = 0
    if m:
        factor = float(m.group(0))
        start = m.end()
    pos = start
    rest = name
    while True:
        stop = len(rest)
        for ch in "*/":
            i = rest.find(ch, pos)
            if i != -1:
                stop = min(stop, i)
        group = rest[pos:stop]
        exponent = 1.0
        if "^" in group:
            group, e = group.split("^", 1)
            exponent = float(e)
        unit = group.strip()
        if pos != 0 and rest[pos - 1] == "/":
            exponent = -exponent
        if unit in _KNOWN:
            factor *= math.pow(_KNOWN[unit], exponent)
        elif unit[:1] in _PREFIX and unit[1:] in _KNOWN:
            factor *= math.pow(_KNOWN[unit[1:]] * _PREFIX[unit[:1]], exponent)
        else:
            raise MechanismError("unknown unit '%s' in unit string '%s'" % (unit, name))
        if stop == len(rest):
            break
        pos = stop + 1
    return factor


def _energy_class(name):
    toks = [t.split("^")[0].strip() for t in re.split(r"[*/]", name)]
    if toks == ["K"]:
        return "K"
    if any(t == "eV" or (t[1:] == "eV" and t[:1] in _PREFIX) for t in toks):
        return "eV"
    return "E"


class UnitSystem:
    """UnitSystem (Units.cpp:390-520): the default units of an input file's `units` block."""

    def __init__(self, block=None):
        self.block = dict(block or {})
        self.length = self.time = self.quantity = self.mass = self.pressure = self.energy = 1.0
        self.activation_energy = 1.0
        explicit_ea = False
        for k, v in (block or {}).items():
            if k == "length":
                self.length = unit_factor(v)
            elif k == "time":
                self.time = unit_factor(v)
            elif k == "quantity":
                self.quantity = unit_factor(v)
            elif k == "mass":
                self.mass = unit_factor(v)
            elif k == "pressure":
                self.pressure = unit_factor(v)
            elif k == "energy":
                self.energy = unit_factor(v)
            elif k == "activation-energy":
                cls = _energy_class(v)
                f = unit_factor(v)
                self.activation_energy = f if cls == "E" else (GAS_CONSTANT * f if cls == "K" else AVOGADRO * f)
                explicit_ea = True
            elif k in ("temperature", "current"):
                pass
            else:
                raise MechanismError("unknown entry '%s' in units block" % k)
        if not explicit_ea:
            self.activation_energy = self.energy / self.quantity

    def overlaid(self, block):
        """The unit system of a node with a local `units` entry: the given defaults replace the inherited ones."""
        return UnitSystem({**self.block, **block}) if block else self

    def rate_coeff(self, v, length_dim, quantity_dim, time_dim=-1.0):
        """convertRateCoeff -> convert -> convertTo (Units.cpp:562-571,618-658) to SI rate-coefficient units of the
        given dimensions; the multiplication order of convertTo is kept."""
        value, units = _split_unit(v)
        if units:
            if not _same_dims(unit_dims(units), (length_dim, 0.0, time_dim, quantity_dim, 0.0)):
                raise MechanismError("incompatible units '%s' for a rate coefficient of dimensions m^%g kmol^%g s^%g"
                                     % (units, length_dim, quantity_dim, time_dim))
            return value * unit_factor(units) / 1.0
        return (value / 1.0 * math.pow(self.mass, 0.0) * math.pow(self.length, length_dim) * math.pow(self.time, time_dim)
                * math.pow(self.quantity, quantity_dim) * math.pow(self.pressure, 0.0) * math.pow(self.energy, 0.0))

    def activation_energy_over_R(self, v):
        """convertActivationEnergy(v, "K") (Units.cpp:676-726,743-758)."""
        value, units = _split_unit(v)
        if not units:
            return value * self.activation_energy / GAS_CONSTANT
        cls, f = _energy_class(units), unit_factor(units)
        want = {"E": (2, 1, -2, -1, 0), "K": (0, 0, 0, 0, 1), "eV": (2, 1, -2, 0, 0)}[cls]
        if not _same_dims(unit_dims(units), want):
            raise MechanismError("don't understand units '%s' as an activation energy" % units)
        value *= f if cls == "E" else (GAS_CONSTANT * f if cls == "K" else AVOGADRO * f)
        return value / (GAS_CONSTANT * 1.0)

    def pressure_pa(self, v):
        value, units = _split_unit(v)
        return value * unit_factor(units) / 1.0 if units else value / 1.0 * self.pressure

    def temperature_k(self, v):
        value, units = _split_unit(v)
        if units not in ("", "K"):
            raise MechanismError("temperature in '%s' is not supported" % units)
        return value


# -------------------------------------------------------------------------------- phase contents

def _find_file(name, dirs):
    for d in dirs:
        p = os.path.join(d, name)
        if os.path.exists(p):
            return p
    raise MechanismError("input file '%s' not found in %s" % (name, list(dirs)))


class _Doc:
    """One input file and the files its sections refer to ("gri30.yaml/species")."""

    def __init__(self, path, dirs):
        self.path = path
        self.dirs = [os.path.dirname(os.path.abspath(path))] + list(dirs)
        self.root = _load(path)
        self.units = UnitSystem(self.root.get("units"))
        self._others = {}

    def section(self, ref):
        """-> (list of entries, UnitSystem of the file that holds them)"""
        if "/" in ref:
            fname, sec = ref.rsplit("/", 1)
            if fname not in self._others:
                self._others[fname] = _Doc(_find_file(fname, self.dirs), self.dirs[1:])
            return self._others[fname].section(sec)
        if ref not in self.root:
            raise MechanismError("section '%s' not found in %s" % (ref, self.path))
        return self.root[ref], self.units


def _phase_species(doc, phase):
    """The species entries of a phase in phase order: `species: all`, a list of names (from the `species`
    section), or a list of {section: all | [names]} maps."""
    spec = phase.get("species", "all")
    out = []

    def take(section, names):
        entries, _ = doc.section(section)
        if names == "all":
            out.extend(entries)
            return
        byname = {str(e["name"]): e for e in entries}
        for n in names:
            if str(n) not in byname:
                raise MechanismError("species '%s' not found in section '%s'" % (n, section))
            out.append(byname[str(n)])

    if spec == "all":
        take("species", "all")
    else:
        plain = [s for s in spec if not isinstance(s, dict)]
        if plain and len(plain) != len(spec):
            raise MechanismError("mixed species list in phase '%s'" % phase.get("name"))
        if plain:
            take("species", plain)
        else:
            for item in spec:
                for section, names in item.items():
                    take(section, names)
    return out


def _phase_reactions(doc, phase):
    """[(entry, UnitSystem, skip_undeclared)] of a phase: `reactions: all | declared-species | none`, a list of
    section names, or a list of {section: all | declared-species} maps; default `all` when the file has a
    `reactions` section."""
    spec = phase.get("reactions")
    if spec is None:
        if "reactions" not in doc.root:
            raise MechanismError("phase '%s' has a kinetics model but the input file has no 'reactions' section "
                                 "(say 'reactions: none' for a phase without reactions)" % phase.get("name"))
        spec = "all"
    out = []

    def take(section, rule):
        if rule == "none":
            return
        if rule not in ("all", "declared-species"):
            raise MechanismError("reaction rule '%s' is not supported" % (rule,))
        entries, units = doc.section(section)
        out.extend((e, units, rule == "declared-species") for e in entries)

    if isinstance(spec, str):
        take("reactions", spec)
    else:
        for item in spec:
            if isinstance(item, dict):
                for section, rule in item.items():
                    take(section, rule)
            else:
                take(item, "all")
    return out


# ------------------------------------------------------------------------------------- reactions

def _parse_equation(eq):
    """parseReactionEquation (Reaction.cpp:889-956): -> (reactants, products, reversible) with the
    pseudo-species "M" / "(+M)" / "(+AR)" still in the maps."""
    tokens = eq.split()
    tokens.append("+")
    reac, prod = {}, {}
    side = reac
    reversible = True
    last_used = None
    for i in range(1, len(tokens)):
        t = tokens[i]
        if t == "+" or t.startswith("(+") or t in ("<=>", "=", "=>"):
            species = tokens[i - 1]
            stoich = 1.0
            gap = i - (last_used if last_used is not None else -1)
            if last_used is not None and tokens[last_used] == "(+" and species.endswith(")"):
                species = "(+" + species         # "(+ M)"
            elif last_used == i - 1 and species.startswith("(+") and species.endswith(")"):
                pass                             # "(+M)"
            elif gap == 2:
                pass                             # species without coefficient
            elif gap == 3:
                stoich = float(tokens[i - 2])
            else:
                raise MechanismError("error parsing reaction equation '%s'" % eq)
            side[species] = side.get(species, 0.0) + stoich
            last_used = i
        if t in ("<=>", "="):
            reversible, side = True, prod
        elif t == "=>":
            reversible, side = False, prod
    return reac, prod, reversible


def _detect_third_body(reac, prod, rate_type, node, eq):
    """Reaction::setEquation (Reaction.cpp:389-555): returns the collider string ("M", "(+M)", "(+AR)", "AR")
    or None, and strips it from both sides."""
    if rate_type == "elementary":
        return None
    third, count, count_m, count_f = "", 0, 0, 0
    for name in sorted(reac):
        if name in prod:
            if count_f == 0:
                third = name
            generic = third in ("M", "(+M)", "(+ M)")
            count += 1
            count_m += generic
            if reac[name] > 1 and prod.get(third, 0.0) > 1:
                count += 1
                count_m += generic
            if name.startswith("(+") and name.endswith(")"):
                count_f += 1
    three_body_type = rate_type.startswith("three-body")
    if count == 0:
        if three_body_type:
            raise MechanismError("reaction '%s' has no valid third-body collider" % eq)
        return None
    if count_f == 1:
        pass
    elif count_m > 1:
        raise MechanismError("multiple generic colliders in '%s'" % eq)
    elif count > 1:
        if count_m:
            third = "M"
        elif "efficiencies" in node:
            effs = node["efficiencies"]
            if len(effs) != 1 or str(next(iter(effs))) not in reac:
                raise MechanismError("ambiguous third-body colliders in '%s'" % eq)
            third = str(next(iter(effs)))
        elif "default-efficiency" in node or three_body_type:
            raise MechanismError("ambiguous third-body colliders in '%s'" % eq)
        else:
            return None
    elif third != "M" and not three_body_type and not third.startswith("(+"):
        # an apparent specific collider: only if the coefficients are integers and one side has three species
        nreac = nprod = 0
        for s in reac.values():
            if math.trunc(s) != s:
                return None
            nreac += int(s)
        for s in prod.values():
            if math.trunc(s) != s:
                return None
            nprod += int(s)
        if nreac != 3 and nprod != 3:
            return None
    for side in (reac, prod):
        if math.trunc(side[third]) != 1:
            side[third] -= 1.0
        else:
            del side[third]
    return third


def _arrhenius(rate, units, conc):
    """ArrheniusBase::setRateParameters (Arrhenius.cpp:35-74) for a rate coefficient whose units are
    (standard concentration)^conc / s (ideal gas: kmol/m^3; conc = 1 - overall order): -> (A, b, Ea/R)."""
    if isinstance(rate, dict):
        A, b, Ea = rate["A"], rate["b"], rate.get("Ea", 0.0)
    else:
        A, b = rate[0], rate[1]
        Ea = rate[2] if len(rate) > 2 else 0.0
    return units.rate_coeff(A, -3.0 * conc, conc), _num(b), units.activation_energy_over_R(Ea)


def _recip_or_inf(c):
    return math.inf if abs(c) < SMALL_NUMBER else 1.0 / c


def _duplicate_ratio(r1, r2):
    """Kinetics::checkDuplicateStoich (Kinetics.cpp:251-293): the multiple c with r2 = c r1 (same direction) or
    r2 = c reversed(r1), 0 when the stoichiometries differ."""
    keys = set(r1) | set(r2)
    k1 = min(r1)
    g1, g2 = (lambda k: r1.get(k, 0.0)), (lambda k: r2.get(k, 0.0))
    if g1(k1) and g2(k1):
        ratio = g2(k1) / g1(k1)
        if not any((g1(k) and not g2(k)) or (not g1(k) and g2(k)) or (g1(k) and abs(g2(k) / g1(k) - ratio) > 1e-8) for k in keys):
            return ratio
    if g1(k1) == 0.0 or g2(-k1) == 0.0:
        return 0.0
    ratio = g2(-k1) / g1(k1)
    if any((g1(k) and not g2(-k)) or (not g1(k) and g2(-k)) or (g1(k) and abs(g2(-k) / g1(k) - ratio) > 1e-8) for k in keys):
        return 0.0
    return ratio


def _check_duplicates(recs, index, names, explicit_tb_rule):
    """Kinetics::checkDuplicates (Kinetics.cpp:112-249): undeclared duplicate reactions and unmatched `duplicate`
    flags are errors; `explicit-third-body-duplicates: modify-efficiency` zeroes the explicit collider's
    efficiency in the generic-M partner (the only way this check changes the tables)."""
    participants = {}
    nets = []
    unmatched = {i for i, r in enumerate(recs) if r["duplicate"]}
    errors = []
    for i, R in enumerate(recs):
        key, net = 0, {}
        for n, st in R["reac"].items():
            k = index[n]
            key += k * (k + 1)
            net[-1 - k] = net.get(-1 - k, 0.0) - st
        for n, st in R["prod"].items():
            k = index[n]
            key += k * (k + 1)
            net[1 + k] = net.get(1 + k, 0.0) + st
        nets.append(net)
        for m in participants.get(key, []):
            O = recs[m]
            if R["duplicate"] and O["duplicate"]:
                unmatched.discard(i); unmatched.discard(m)
                continue
            if R["type"] != O["type"]:
                continue
            c = _duplicate_ratio(nets[i], nets[m])
            if c == 0:
                continue
            if c < 0.0 and not R["reversible"] and not O["reversible"]:
                continue
            if R["tb"] is not None and O["tb"] is not None:
                t1, t2 = R["tb"], O["tb"]
                eff = lambda t, sp: t["effs"].get(sp, t["default"])  # noqa: E731
                if not any(eff(t1, sp) * eff(t2, sp) != 0.0 for sp in names):
                    continue
                if (t1["name"] == "M") != (t2["name"] == "M"):
                    if explicit_tb_rule == "mark-duplicate":
                        R["duplicate"] = O["duplicate"] = True
                        continue
                    if explicit_tb_rule == "modify-efficiency":
                        if t1["name"] == "M":
                            t1["effs"][t2["name"]] = 0.0
                        else:
                            t2["effs"][t1["name"]] = 0.0
                        continue
                    if explicit_tb_rule == "warn":
                        continue
            errors.append("undeclared duplicate reactions: '%s' and '%s'" % (O["eq"], R["eq"]))
        participants.setdefault(key, []).append(i)
    for i in sorted(unmatched):
        errors.append("no duplicate found for the declared duplicate reaction '%s'" % recs[i]["eq"])
    if errors:
        raise MechanismError("; ".join(errors))


def tables_from_yaml(path, phase="", data_dirs=()):
    """Flat mechanism tables (the keys of a .ctbm file, include/ctb_clib.h ctb_mech_tables) of phase `phase`
    (default: the first phase) of a YAML input file in the reference's format."""
    doc = _Doc(path, data_dirs)
    phases = doc.root.get("phases") or []
    ph = next((p for p in phases if not phase or str(p.get("name")) == phase), None)
    if ph is None:
        raise MechanismError("phase '%s' not found in %s" % (phase, path))
    if ph.get("thermo") != "ideal-gas":
        raise MechanismError("only the ideal-gas phase model is on the batched path; got '%s'" % ph.get("thermo"))
    if ph.get("kinetics") not in ("gas", "bulk"):
        raise MechanismError("only bulk (gas) kinetics is on the batched path; got '%s'" % ph.get("kinetics"))

    # ---- species: NASA7 thermo and molecular weights (Phase::addSpecies sums in PHASE element order)
    species = _phase_species(doc, ph)
    names = [str(s["name"]) for s in species]
    index = {n: k for k, n in enumerate(names)}
    if len(index) != len(names):
        raise MechanismError("duplicate species in phase '%s'" % ph.get("name"))
    K = len(names)
    # elements (setupPhase, ThermoFactory.cpp): a declared list is closed (species with other elements are an
    # error, or dropped under skip-undeclared-elements); without a list the elements are added as met
    declared = "elements" in ph
    elements = [str(e) for e in ph.get("elements", [])]
    skip_elements = bool(ph.get("skip-undeclared-elements", False))
    kept = []
    for s in species:
        missing = [e for e in sorted(str(x) for x in s.get("composition", {})) if e not in elements]
        if missing and declared:
            if skip_elements:
                continue
            raise MechanismError("species '%s' contains the undeclared element(s) %s" % (s["name"], missing))
        elements += missing
        kept.append(s)
    species = kept
    names = [str(s["name"]) for s in species]
    index = {n: k for k, n in enumerate(names)}
    K = len(names)
    molwt, tmid, tmin, tmax, low, high = [], [], [], [], [], []
    p_ref = None
    for s in species:
        comp = {str(e): float(n) for e, n in s.get("composition", {}).items()}
        wt = 0.0
        for e in elements:
            if e not in ATOMIC_WEIGHTS:
                raise MechanismError("no atomic weight for element '%s'" % e)
            wt += comp.get(e, 0.0) * ATOMIC_WEIGHTS[e]
        molwt.append(max(wt, 1e-20))
        th = s.get("thermo", {})
        if th.get("model") != "NASA7":
            raise MechanismError("species '%s' uses thermo model '%s'; only NASA7 is on the batched path"
                                 % (s["name"], th.get("model")))
        tr = [doc.units.temperature_k(x) for x in th["temperature-ranges"]]
        data = [[_num(x) for x in row] for row in th["data"]]
        if len(tr) not in (2, 3) or len(data) != len(tr) - 1 or any(len(r) != 7 for r in data):
            raise MechanismError("species '%s': malformed NASA7 data" % s["name"])
        pr = doc.units.pressure_pa(th["reference-pressure"]) if "reference-pressure" in th else ONE_ATM
        p_ref = pr if p_ref is None else p_ref
        if pr != p_ref:
            raise MechanismError("species '%s' has a different reference pressure" % s["name"])
        tmin.append(tr[0]); tmax.append(tr[-1]); tmid.append(tr[1])
        low.append(data[0]); high.append(data[1] if len(tr) == 3 else data[0])

    # ---- reactions
    T = {k: [] for k in ("rxn_type", "rxn_flags", "rate_A", "rate_b", "rate_Ea_R", "low_A", "low_b", "low_Ea_R",
                         "falloff_coeffs", "multiplier", "delta_n", "reac_species", "reac_stoich", "reac_order",
                         "prod_species", "prod_stoich", "tb_default", "tb_species", "tb_eff", "plog_P", "plog_A",
                         "plog_b", "plog_Ea_R", "cheb_nT", "cheb_nP", "cheb_limits", "cheb_coeffs")}
    ptr = {k: [0] for k in ("reac_ptr", "prod_ptr", "tb_ptr", "plog_ptr", "cheb_ptr")}
    equations = []
    recs = []
    skip_third_bodies = bool(ph.get("skip-undeclared-third-bodies", False))
    for node, units, skip_undeclared in _phase_reactions(doc, ph):
        eq = str(node["equation"])
        units = units.overlaid(node.get("units"))
        rtype = str(node.get("type", "Arrhenius"))
        reac, prod, reversible = _parse_equation(eq)
        third = _detect_third_body(reac, prod, rtype, node, eq)
        orders = {str(k): _num(v) for k, v in (node.get("orders") or {}).items()}
        undeclared = [n for n in list(reac) + list(prod) + list(orders) if n not in index]
        # a specific collider written as a species ("+ AR") takes part in the parse-time validity check
        # (Reaction.cpp:932-936); one written "(+AR)" is caught by ThirdBody::checkSpecies (Reaction.cpp:858-882)
        collider = None if third is None else (third[2:-1].strip() if third.startswith("(+") else third)
        if collider not in (None, "M") and collider not in index:
            if not third.startswith("(+"):
                undeclared.append(collider)
            elif skip_third_bodies:
                continue
            else:
                raise MechanismError("reaction '%s' has the undeclared third body '%s'" % (eq, collider))
        if undeclared:
            if skip_undeclared:
                continue
            raise MechanismError("reaction '%s' contains undeclared species %s" % (eq, undeclared))
        if orders and reversible:
            raise MechanismError("reaction orders may only be given for irreversible reactions ('%s')" % eq)
        flags = RXN_REVERSIBLE if reversible else 0
        kind = rtype[11:] if rtype.startswith("three-body-") else rtype
        if kind in ("three-body", "elementary"):
            kind = "Arrhenius"
        falloff = kind in ("falloff", "chemically-activated")

        # third body (Reaction.cpp:777-830, BulkKinetics.cpp:66-84)
        tb_default, effs, mass_action = 0.0, {}, False
        if third is not None:
            mass_action = not third.startswith("(+")
            tname = third[2:-1].strip() if third.startswith("(+") else third
            if tname == "M":
                tb_default = _num(node.get("default-efficiency", 1.0))
                effs = {str(k): _num(v) for k, v in (node.get("efficiencies") or {}).items()}
            else:
                if _num(node.get("default-efficiency", 0.0)) != 0.0:
                    raise MechanismError("invalid default efficiency for the explicit collider of '%s'" % eq)
                effs = {str(k): _num(v) for k, v in (node.get("efficiencies") or {tname: 1.0}).items()}
                if list(effs) != [tname]:
                    raise MechanismError("incompatible third-body definitions in '%s'" % eq)
            if falloff:
                if mass_action:
                    raise MechanismError("third-body collider of '%s' does not use the '(+M)' notation" % eq)
            elif kind == "Chebyshev" and tname == "M":
                third = None  # deprecated notation: the collider is dropped (Reaction.cpp:189-194)
            elif kind == "pressure-dependent-Arrhenius" and tname == "M":
                raise MechanismError("superfluous 'M' in pressure-dependent-Arrhenius reaction '%s'" % eq)
        elif falloff:
            raise MechanismError("falloff reaction '%s' has no pressure-dependent third body" % eq)
        elif "efficiencies" in node or "default-efficiency" in node:
            raise MechanismError("reaction '%s' specifies efficiencies but has no third body" % eq)
        if third is not None:
            flags |= RXN_THIRD_BODY | (RXN_MASS_ACTION if mass_action else 0)
            T["tb_default"].append(tb_default)
            unknown = [n for n in effs if n not in index]
            if unknown and not skip_third_bodies:
                raise MechanismError("reaction '%s' defines third-body efficiencies for undeclared species %s" % (eq, unknown))
        else:
            T["tb_default"].append(0.0)

        # stoichiometry in the order of the reference's maps (std::map: by name), orders appended (Kinetics.cpp:734-777)
        rk = [index[n] for n in sorted(reac)]
        rst = [reac[n] for n in sorted(reac)]
        rord = list(rst)
        for n in sorted(orders):
            if index[n] in rk:
                rord[rk.index(index[n])] = orders[n]
            else:
                rk.append(index[n]); rst.append(0.0); rord.append(orders[n])
        T["reac_species"] += rk; T["reac_stoich"] += rst; T["reac_order"] += rord
        T["prod_species"] += [index[n] for n in sorted(prod)]
        T["prod_stoich"] += [prod[n] for n in sorted(prod)]
        ptr["reac_ptr"].append(len(T["reac_species"])); ptr["prod_ptr"].append(len(T["prod_species"]))
        dn = 0.0
        for n in sorted(prod):
            dn += prod[n]
        for n in sorted(reac):
            dn -= reac[n]
        T["delta_n"].append(dn)

        # exponent of the standard concentration in the units of the rate coefficient, accumulated in the
        # order of Reaction::calculateRateCoeffUnits (Reaction.cpp:576-617): orders override stoichiometry
        conc = 1.0
        for n in sorted(orders):
            conc -= orders[n]
        for n in sorted(reac):
            if n not in orders:
                conc -= reac[n]
        if third is not None and mass_action:
            conc -= 1.0

        A = b = EaR = lA = lb = lEaR = 0.0
        fc = [0.0] * 5
        cheb = [0.0] * 4
        nT = nP = 0
        if kind == "Arrhenius":
            code = RATE_ARRHENIUS
            A, b, EaR = _arrhenius(node["rate-constant"], units, conc)
        elif falloff:
            chem = kind == "chemically-activated"
            if chem:
                flags |= RXN_CHEM_ACTIVATED
            # Falloff.cpp:130-161: the low-pressure limit carries one more concentration (falloff), the
            # high-pressure limit one less (chemically activated)
            A, b, EaR = _arrhenius(node["high-P-rate-constant"], units, conc + 1.0 if chem else conc)
            lA, lb, lEaR = _arrhenius(node["low-P-rate-constant"], units, conc if chem else conc - 1.0)
            if "Troe" in node:
                code, f = RATE_TROE, node["Troe"]
                fc[0] = _num(f["A"]); fc[1] = _recip_or_inf(_num(f["T3"])); fc[2] = _recip_or_inf(_num(f["T1"]))
                fc[3] = _num(f["T2"]) if "T2" in f else 0.0
            elif "SRI" in node:
                code, f = RATE_SRI, node["SRI"]
                fc[0], fc[1], fc[2] = _num(f["A"]), _num(f["B"]), _num(f["C"])
                full = "D" in f and "E" in f
                fc[3], fc[4] = (_num(f["D"]), _num(f["E"])) if full else (1.0, 0.0)
                if ("D" in f) != ("E" in f):
                    raise MechanismError("SRI falloff of '%s' needs 3 or 5 parameters" % eq)
            elif "Tsang" in node:
                code, f = RATE_TSANG, node["Tsang"]
                fc[0], fc[1] = _num(f["A"]), _num(f.get("B", 0.0))
            else:
                code = RATE_LINDEMANN
        elif kind == "pressure-dependent-Arrhenius":
            code = RATE_PLOG
            entries = [(units.pressure_pa(r["P"]), i, _arrhenius(r, units, conc)) for i, r in enumerate(node["rate-constants"])]
            # PlogRate::validate (PlogRate.cpp:125-158): the summed rate at every pressure must be positive
            for P0 in sorted({e[0] for e in entries}):
                for Tv in (300.0, 500.0, 1000.0, 2000.0, 5000.0, 10000.0):
                    ksum = sum(pa * math.exp(pb * math.log(Tv) - pe / Tv) for P, _, (pa, pb, pe) in entries if P == P0)
                    if not ksum > 0:
                        raise MechanismError("invalid rate coefficient for reaction '%s' at P = %.5g, T = %.1f" % (eq, P0, Tv))
            # std::multimap: by P, stable; PlogRate keeps log(P) and reports exp(log(P)) (PlogRate.cpp:92-123)
            for P, _, (pa, pb, pe) in sorted(entries, key=lambda x: (x[0], x[1])):
                T["plog_P"].append(math.exp(math.log(P))); T["plog_A"].append(pa); T["plog_b"].append(pb); T["plog_Ea_R"].append(pe)
        elif kind == "Chebyshev":
            code = RATE_CHEBYSHEV
            data = [[_num(x) for x in row] for row in node["data"]]
            nT, nP = len(data), len(data[0])
            if any(len(r) != nP for r in data):
                raise MechanismError("inconsistent Chebyshev coefficient matrix in '%s'" % eq)
            data[0][0] += math.log10(units.rate_coeff(1.0, -3.0 * conc, conc))  # ChebyshevRate.cpp:82-83
            cheb = [units.temperature_k(node["temperature-range"][0]), units.temperature_k(node["temperature-range"][1]),
                    units.pressure_pa(node["pressure-range"][0]), units.pressure_pa(node["pressure-range"][1])]
            for row in data:
                T["cheb_coeffs"] += row
        else:
            raise MechanismError("rate type '%s' of reaction '%s' is not on the batched path" % (rtype, eq))
        ptr["plog_ptr"].append(len(T["plog_P"])); ptr["cheb_ptr"].append(len(T["cheb_coeffs"]))
        T["rxn_type"].append(code); T["rxn_flags"].append(flags)
        T["rate_A"].append(A); T["rate_b"].append(b); T["rate_Ea_R"].append(EaR)
        T["low_A"].append(lA); T["low_b"].append(lb); T["low_Ea_R"].append(lEaR)
        T["falloff_coeffs"] += fc; T["multiplier"].append(1.0)
        T["cheb_nT"].append(nT); T["cheb_nP"].append(nP); T["cheb_limits"] += cheb
        # Reaction::check / ArrheniusBase::check (Reaction.cpp:137-165, Arrhenius.cpp:113-124)
        if not node.get("nonreactant-orders", False) and any(n not in reac for n in orders):
            raise MechanismError("reaction order specified for a non-reactant species in '%s'" % eq)
        if not node.get("negative-orders", False) and any(v < 0.0 for v in orders.values()):
            raise MechanismError("negative reaction order in '%s'" % eq)
        if not node.get("negative-A", False) and (A < 0.0 or lA < 0.0):
            raise MechanismError("negative pre-exponential factor in '%s' (needs negative-A: true)" % eq)
        recs.append({"reac": dict(reac), "prod": dict(prod), "reversible": reversible, "duplicate": bool(node.get("duplicate", False)),
                     "type": (code, flags & (RXN_THIRD_BODY | RXN_MASS_ACTION | RXN_CHEM_ACTIVATED)), "eq": eq,
                     "tb": None if third is None else {"name": tname, "default": tb_default, "effs": effs}})
        # Reaction::equation (Reaction.cpp:344-387): sides in map order, the collider appended
        collider = "" if third is None else ((" + " + tname) if mass_action else (" (+" + tname + ")"))

        def side_string(side):
            return " + ".join(("%g " % side[n] if side[n] != 1.0 else "") + n for n in sorted(side)) + collider
        equations.append(side_string(reac) + (" <=> " if reversible else " => ") + side_string(prod))

    _check_duplicates(recs, index, names, str(ph.get("explicit-third-body-duplicates", "warn")))
    for rec in recs:  # third-body efficiencies of declared species, by species index (BulkKinetics.cpp:66-84)
        if rec["tb"] is not None:
            effs = rec["tb"]["effs"]
            for k in sorted(index[n] for n in effs if n in index):
                T["tb_species"].append(k)
                T["tb_eff"].append(effs[names[k]])
        ptr["tb_ptr"].append(len(T["tb_species"]))

    ints = ("rxn_type", "rxn_flags", "reac_species", "prod_species", "tb_species", "cheb_nT", "cheb_nP")
    out = {"dims": np.array([K, len(equations)], dtype=np.int32), "p_ref": np.array([p_ref if p_ref is not None else ONE_ATM]),
           "mol_weights": np.array(molwt), "nasa_tmid": np.array(tmid), "nasa_tmin": np.array(tmin), "nasa_tmax": np.array(tmax),
           "nasa_low": np.array(low, dtype=np.float64).reshape(-1), "nasa_high": np.array(high, dtype=np.float64).reshape(-1)}
    order_keys = ["rxn_type", "rxn_flags", "rate_A", "rate_b", "rate_Ea_R", "low_A", "low_b", "low_Ea_R", "falloff_coeffs",
                  "multiplier", "delta_n", "reac_ptr", "reac_species", "reac_stoich", "reac_order", "prod_ptr", "prod_species",
                  "prod_stoich", "tb_default", "tb_ptr", "tb_species", "tb_eff", "plog_ptr", "plog_P", "plog_A", "plog_b",
                  "plog_Ea_R", "cheb_ptr", "cheb_nT", "cheb_nP", "cheb_limits", "cheb_coeffs"]
    for k in order_keys:
        v = ptr[k] if k in ptr else T[k]
        out[k] = np.array(v, dtype=np.int32 if (k in ints or k in ptr) else np.float64)
    out["species_names"] = "\n".join(names)
    out["equations"] = "\n".join(equations)
    return out
