"""Bmad lattice files to a ``Segment`` (converters/bmad.py:17-693).

Host logic only; a lattice data format in front of the tracking path.  The file is a small
program: variables, element definitions with arithmetic in their attributes, attribute
assignments (``q1[k1] = 2*q2[k1]``, wildcards in the name), beam lines and ``use``.  It is
executed statement by statement into a namespace; expressions are parsed with ``ast`` and
evaluated over a closed set of operators, functions and constants (nothing is ``eval``-ed).

Covered: comments, ``&`` and trailing-``,`` continuation, ``call, file = ...`` (with
``$ENV`` path parts), inheritance from a previously defined element, nested lines, ``n*name``
repetition and ``-name`` reversal in lines.  Overlays / groups are parsed and skipped (they are
control-room knobs, not elements).  Element kinds the model does not have become drifts of the
same length, as in the reference (``converters/bmad.py:627-633``).
"""

import ast
import fnmatch
import math
import operator
import os
import re
from pathlib import Path

import torch
from scipy import constants as _c

_MC2_EV = _c.physical_constants["electron mass energy equivalent in MeV"][0] * 1e6
CONSTANTS = {  # converters/bmad.py:381-394
    "pi": math.pi, "twopi": 2 * math.pi, "c_light": _c.c, "emass": _MC2_EV * 1e-9,
    "m_electron": _MC2_EV, "raddeg": _c.degree,
}
FUNCTIONS = {"sqrt": math.sqrt, "asin": math.asin, "sin": math.sin, "cos": math.cos,
             "tan": math.tan, "acos": math.acos, "atan": math.atan, "exp": math.exp,
             "log": math.log, "abs": abs, "abs_func": abs}
_BINARY = {ast.Add: operator.add, ast.Sub: operator.sub, ast.Mult: operator.mul,
           ast.Div: operator.truediv, ast.Pow: operator.pow}


class BmadError(ValueError):
    pass


# ------------------------------------------------------------------ expressions
def evaluate(expression: str, names: dict):
    """Number for an arithmetic expression over variables, constants and ``elem[attr]``
    references; anything that is not arithmetic (``open``, ``electron``, quoted text) is
    returned as a string."""
    text = expression.strip()
    if len(text) >= 2 and text[0] in "\"'" and text[-1] == text[0]:
        return text[1:-1]
    refs = {}

    def lift(match):  # elem[attr] -> placeholder name
        key = f"__ref{len(refs)}__"
        refs[key] = (match.group(1), match.group(2))
        return key

    py = re.sub(r"([a-z_][a-z0-9_.]*)\[([a-z0-9_%]+)\]", lift, text).replace("^", "**")
    try:
        tree = ast.parse(py, mode="eval")
        return _walk(tree.body, names, refs)
    except (SyntaxError, BmadError, KeyError, TypeError, ArithmeticError, ValueError):
        return text


def _walk(node, names, refs):
    if isinstance(node, ast.Constant) and isinstance(node.value, (int, float)):
        return node.value
    if isinstance(node, ast.BinOp) and type(node.op) in _BINARY:
        return _BINARY[type(node.op)](_walk(node.left, names, refs), _walk(node.right, names, refs))
    if isinstance(node, ast.UnaryOp) and isinstance(node.op, (ast.USub, ast.UAdd)):
        value = _walk(node.operand, names, refs)
        return -value if isinstance(node.op, ast.USub) else value
    if isinstance(node, ast.Call) and isinstance(node.func, ast.Name) and node.func.id in FUNCTIONS:
        return FUNCTIONS[node.func.id](*[_walk(a, names, refs) for a in node.args])
    if isinstance(node, ast.Name):
        if node.id in refs:
            element, attribute = refs[node.id]
            return names[element][attribute]
        if node.id in CONSTANTS:
            return CONSTANTS[node.id]
        value = names[node.id]
        if isinstance(value, (int, float)):
            return value
    raise BmadError(f"not an arithmetic expression: {ast.dump(node)}")


# ------------------------------------------------------------------ statements
def _expand_path(text: str, base: Path) -> Path:
    path = Path(*[os.environ[p[1:]] if p.startswith("$") else p for p in Path(text).parts])
    return path if path.is_absolute() else base / path


def statements(path: Path) -> list:
    """Comment-free, lower-cased, continuation-merged statements of a file and of the files it
    calls (converters/bmad.py:17-98)."""
    out, pending = [], ""
    for raw in Path(path).read_text().splitlines():
        line = raw.split("!", 1)[0].strip()
        if not line:
            continue
        line = pending + line
        pending = ""
        if line.endswith("&"):
            pending = line[:-1]
            continue
        if line.endswith((",", "{", "(")) or line.count("(") > line.count(")") \
                or line.count("{") > line.count("}"):
            pending = line
            continue
        call = re.fullmatch(r"call\s*,\s*file\s*=\s*(\S+)", line, flags=re.IGNORECASE)
        if call:  # paths and $ENV names are case sensitive: expand before lower-casing
            out.extend(statements(_expand_path(call.group(1).strip("\"'"), Path(path).parent)))
        else:
            out.append(line.lower())
    if pending:
        out.append(pending.lower())
    return out


def _split_top(text: str) -> list:
    """Split at commas that are not inside brackets or quotes."""
    parts, depth, quote, cur = [], 0, "", ""
    for ch in text:
        if quote:
            quote = "" if ch == quote else quote
        elif ch in "\"'":
            quote = ch
        elif ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        elif ch == "," and depth == 0:
            parts.append(cur.strip())
            cur = ""
            continue
        cur += ch
    if cur.strip():
        parts.append(cur.strip())
    return parts


_NAME = r"[a-z_][a-z0-9_.]*"


def parse(path) -> dict:
    """Execute the file into a namespace: numbers for variables, ``{"element_type": ..,
    attr: ..}`` for elements, lists of member names for lines, ``"__use__"`` for the line in
    use (converters/bmad.py:366-411)."""
    names: dict = {}
    for stmt in statements(Path(path)):
        m = re.fullmatch(rf"use\s*,\s*({_NAME})", stmt)
        if m:
            names["__use__"] = m.group(1)
            continue
        m = re.fullmatch(rf"({_NAME})\s*:\s*line\s*=\s*\((.*)\)", stmt)
        if m:
            names[m.group(1)] = _line_members(m.group(2))
            continue
        m = re.fullmatch(rf"({_NAME})\s*:\s*(overlay|group)\s*=.*", stmt)
        if m:
            continue
        m = re.fullmatch(rf"([a-z0-9_.*%]+)\[([a-z0-9_%]+)\]\s*=\s*(.+)", stmt)
        if m:  # attribute assignment, the name may hold wildcards
            pattern, attribute, value = m.group(1), m.group(2), evaluate(m.group(3), names)
            targets = [pattern] if not any(c in pattern for c in "*%") else [
                k for k, v in names.items()
                if isinstance(v, dict) and fnmatch.fnmatchcase(k, pattern.replace("%", "?"))]
            for target in targets:
                names.setdefault(target, {})[attribute] = value
            continue
        m = re.fullmatch(rf"({_NAME})\s*:\s*({_NAME})\s*(?:,\s*(.*))?", stmt)
        if m:  # element definition, possibly inheriting from an earlier element
            name, kind, rest = m.group(1), m.group(2), m.group(3)
            parent = names.get(kind)
            element = dict(parent) if isinstance(parent, dict) and "element_type" in parent \
                else {"element_type": kind}
            for item in _split_top(rest or ""):
                key, _, value = item.partition("=")
                element[key.strip()] = evaluate(value, names) if value else True
            names[name] = element
            continue
        m = re.fullmatch(rf"({_NAME})\s*=\s*(.+)", stmt)
        if m:
            names[m.group(1)] = evaluate(m.group(2), names)
            continue
        # anything else (e.g. "no_digested") has no bearing on the element list
    return names


def _line_members(text: str) -> list:
    members = []
    for item in _split_top(text):
        m = re.fullmatch(rf"(?:(\d+)\s*\*\s*)?(-?)\s*({_NAME})", item)
        if not m:
            raise BmadError(f"cannot read line member {item!r}")
        count, reverse, name = int(m.group(1) or 1), m.group(2) == "-", m.group(3)
        members.extend([("-" + name) if reverse else name] * count)
    return members


# ------------------------------------------------------------------ elements
def _t(value):
    return torch.tensor([float(value)], dtype=torch.float32)


def _check(kind, props, understood):
    """Attributes the conversion does not know must not vanish silently
    (converters/bmad.py:414-431)."""
    for key in props:
        assert key in understood or key == "element_type", (
            f"Property {key} with value {props[key]} for element type {kind} is currently not"
            f" understood. Other values in properties are {list(props)}.")


_COMMON = ("alias", "type", "descrip")


def convert(name: str, names: dict):
    """Element or (nested) ``Segment`` for ``name`` (converters/bmad.py:434-639)."""
    import lynx_b200 as lx

    reverse = name.startswith("-")
    key = name[1:] if reverse else name
    parsed = names[key]
    if isinstance(parsed, list):
        members = [convert(m, names) for m in parsed]
        return lx.Segment(elements=members[::-1] if reverse else members, name=key)
    if not (isinstance(parsed, dict) and "element_type" in parsed):
        raise BmadError(f"Unknown Bmad element type for name = {key!r}")
    kind, p = parsed["element_type"], parsed
    length = _t(p.get("l", 0.0))
    if kind == "marker":
        _check(kind, p, _COMMON + ("sr_wake", "sr_wake%scale_with_length", "sr_wake%amp_scale"))
        return lx.Marker(name=key)
    if kind in ("monitor", "instrument"):
        _check(kind, p, _COMMON + ("l",))
        return lx.Drift(length=length, name=key) if "l" in p else lx.Marker(name=key)
    if kind in ("pipe", "drift"):
        _check(kind, p, _COMMON + ("l",))
        return lx.Drift(length=_t(p["l"]), name=key)
    if kind in ("hkicker", "vkicker"):
        _check(kind, p, _COMMON + ("l", "kick"))
        cls = lx.HorizontalCorrector if kind == "hkicker" else lx.VerticalCorrector
        return cls(length=length, angle=_t(p.get("kick", 0.0)), name=key)
    if kind == "sbend":
        _check(kind, p, _COMMON + ("hgap", "l", "angle", "e1", "e2", "fint", "fintx",
                                   "fringe_type", "ref_tilt", "g", "dg"))
        return lx.Dipole(
            length=_t(p["l"]), gap=_t(p.get("hgap", 0.0)), angle=_t(p.get("angle", 0.0)),
            e1=_t(p.get("e1", 0.0)), e2=_t(p.get("e2", 0.0)), tilt=_t(p.get("ref_tilt", 0.0)),
            fringe_integral=_t(p.get("fint", 0.0)),
            fringe_integral_exit=_t(p["fintx"]) if "fintx" in p else None, name=key)
    if kind == "quadrupole":
        _check(kind, p, _COMMON + ("l", "k1", "aperture", "tilt"))
        return lx.Quadrupole(length=_t(p["l"]), k1=_t(p.get("k1", 0.0)),
                             tilt=_t(p.get("tilt", 0.0)), name=key)
    if kind == "solenoid":
        _check(kind, p, _COMMON + ("l", "ks"))
        return lx.Solenoid(length=_t(p["l"]), k=_t(p.get("ks", 0.0)), name=key)
    if kind == "lcavity":
        _check(kind, p, _COMMON + ("l", "rf_frequency", "voltage", "phi0", "sr_wake",
                                   "cavity_type"))
        # Bmad counts the phase in turns from the crest, with the opposite sign (:578-580)
        return lx.Cavity(length=_t(p["l"]), voltage=_t(p.get("voltage", 0.0)),
                         phase=_t(-math.degrees(p.get("phi0", 0.0) * 2 * math.pi)),
                         frequency=_t(p["rf_frequency"]), name=key)
    if kind in ("rcollimator", "ecollimator"):
        _check(kind, p, _COMMON + ("l", "x_limit", "y_limit"))
        return lx.Aperture(x_max=_t(p.get("x_limit", math.inf)), y_max=_t(p.get("y_limit", math.inf)),
                           shape="rectangular" if kind == "rcollimator" else "elliptical", name=key)
    if kind == "wiggler":
        _check(kind, p, _COMMON + ("l_period", "n_period", "b_max", "l", "tilt", "ds_step"))
        return lx.Undulator(length=_t(p["l"]), name=key)
    if kind == "patch":
        _check(kind, p, ("tilt", "l"))
        return lx.Drift(length=length, name=key)
    import warnings

    warnings.warn(f"Element {key} of type {kind} cannot be converted correctly. Using drift"
                  " section instead.")
    return lx.Drift(length=length, name=key)


def read_bmad_lattice(path, environment_variables=None):
    """``Segment`` of the line named by ``use`` (converters/bmad.py:642-693)."""
    for k, v in (environment_variables or {}).items():
        os.environ[k] = v
    names = parse(_expand_path(str(path), Path.cwd()))
    if "__use__" not in names:
        raise BmadError("the lattice file names no line to use")
    return convert(names["__use__"], names)
