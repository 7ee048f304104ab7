"""NEC deck → wire geometry: a restatement of the reference's `nec-file` crate and of how cem-app turns a
parsed deck into scene colliders, so BASELINE configs[4] ("dipole antenna from a nec-file NEC deck rasterized into the
material lattice") can be expressed the way the application expresses it.

Follows, quirks included (each flagged where it happens):
  * tokenizer / card parser   nec-file/src/parser.rs:84-292, 300-376
  * card vocabulary           nec-file/src/card.rs:13-72 ("SC" parses as SP, card.rs:63), 136-298, 300-395
  * geometry interpreter      nec-file/src/interpreter.rs:113-380 (GW, GM, GR, GS, GX, GE)
  * deck → colliders          cem-app/src/composer/file_formats/nec.rs:38-82

All arithmetic the reference does in f32 is done in numpy float32 here.  Host-side input handling only; nothing in this
module touches the device.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Iterable, Iterator, Optional

import numpy as np

F = np.float32

CARD_TYPES = {"CM": "CM", "CE": "CE", "GA": "GA", "GC": "GC", "GE": "GE", "GM": "GM", "GR": "GR", "GS": "GS", "GW": "GW",
              "GX": "GX", "SP": "SP", "SC": "SP"}   # card.rs:50-64 — "SC" maps to Sp in the reference


class NecError(ValueError):
    """parser.rs `Error` / interpreter.rs `Error`."""


class UnexpectedEnd(NecError):
    pass


class InvalidParameter(NecError):
    pass


class UnexpectedCardForSection(NecError):
    pass


# ---- tokenizer (parser.rs:300-376) ---------------------------------------------------------------------------------
class _Tokens:
    def __init__(self, line: str, section: str):
        self.line, self.pos, self.section = line, 0, section

    def remainder(self) -> str:
        return self.line[self.pos:]

    def token(self) -> Optional[str]:
        n = len(self.line)
        i = self.pos
        while i < n and self.line[i].isspace():
            i += 1
        if i >= n:
            self.pos = n
            return None
        j = i
        while j < n and not self.line[j].isspace():
            j += 1
        self.pos = j
        return self.line[i:j]

    def _need(self) -> str:
        t = self.token()
        if t is None:
            raise UnexpectedEnd(f"unexpected end of card in the {self.section} section")
        return t

    def u32(self) -> int:
        t = self._need()
        # Rust `u32::from_str`: optional '+', decimal digits only
        body = t[1:] if t.startswith("+") else t
        if not body.isdigit() or not body.isascii() or int(body) > 0xFFFFFFFF:
            raise InvalidParameter(t)
        return int(body)

    def f32(self) -> np.float32:
        t = self._need()
        try:
            if "_" in t:
                raise ValueError
            return F(float(t))
        except ValueError:
            raise InvalidParameter(t) from None

    def f32s(self, n: int) -> np.ndarray:
        return np.array([self.f32() for _ in range(n)], dtype=F)

    def card_type(self) -> str:
        t = self._need()
        if t not in CARD_TYPES:
            raise InvalidParameter(t)
        return CARD_TYPES[t]


# ---- geometry records (interpreter.rs:56-98, card.rs:136-151) --------------------------------------------------------
@dataclass
class WireSegments:
    """card.rs:136-146; `kind` = "fixed" (radius) | "tapered" (length_ratio, first_radius, last_radius)."""
    kind: str
    radius: np.float32 = F(0)
    length_ratio: np.float32 = F(1)
    first_radius: np.float32 = F(0)
    last_radius: np.float32 = F(0)

    def scale(self, s) -> None:  # card.rs:185-199
        if self.kind == "fixed":
            self.radius = F(self.radius * s)
        else:
            self.first_radius = F(self.first_radius * s)
            self.last_radius = F(self.last_radius * s)

    def dimensions(self, num_segments: int, total_length) -> Iterator[tuple]:
        """card.rs:149-183, 221-268.  Yields ("flat", length, radius) or ("tapered", length, start_radius, end_radius)."""
        ns = F(num_segments)
        total_length = F(total_length)
        if self.kind == "fixed":
            seg = F(total_length / ns)
            for _ in range(num_segments):
                yield ("flat", seg, self.radius)
            return
        radius_ratio = F(np.power(F(self.last_radius / self.first_radius), F(F(1) / F(ns - F(1)))))
        if self.length_ratio == F(1):
            length = F(total_length / ns)
        else:
            length = F(F(total_length * F(F(1) - self.length_ratio)) / F(F(1) - F(np.power(self.length_ratio, ns))))
        start = self.first_radius
        for _ in range(num_segments):
            nxt = F(start * radius_ratio)
            yield ("tapered", length, start, nxt)
            length = F(length * self.length_ratio)
            start = nxt

    def copy(self) -> "WireSegments":
        return WireSegments(self.kind, self.radius, self.length_ratio, self.first_radius, self.last_radius)


@dataclass
class Geometry:
    """interpreter.rs:56-66: a specification + a homogeneous 4×4 f32 transform."""
    kind: str                              # "wire" | "wire_arc"
    transform: np.ndarray                  # 4×4 float32
    length: np.float32 = F(0)
    num_segments: int = 0
    segments: Optional[WireSegments] = None
    arc_radius: np.float32 = F(0)
    arc_angles: tuple = (F(0), F(0))
    wire_radius: np.float32 = F(0)

    def copy(self) -> "Geometry":
        return Geometry(self.kind, self.transform.copy(), self.length, self.num_segments,
                        None if self.segments is None else self.segments.copy(), self.arc_radius, self.arc_angles, self.wire_radius)

    def append_transform(self, m: np.ndarray) -> None:   # interpreter.rs:62-66
        self.transform = _matmul4(m, self.transform)


@dataclass
class NecFile:
    """interpreter.rs:37-45."""
    comments: list = field(default_factory=list)
    geometry: list = field(default_factory=list)       # [(tag, Geometry)] sorted by tag (BTreeMap order)
    ground_plane: Optional[bool] = None                # None = not present; True/False = current_modified
    symmetry: tuple = ("none",)
    ignored_decks: list = field(default_factory=list)

    @staticmethod
    def from_lines(lines: Iterable[str]) -> "NecFile":
        return parse_deck(lines)

    @staticmethod
    def from_text(text: str) -> "NecFile":
        return parse_deck(text.splitlines())


# ---- small f32 linear algebra (nalgebra Matrix4<f32> / Isometry3<f32> as the interpreter uses them) ----------------
def _matmul4(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.zeros((4, 4), dtype=F)
    for r in range(4):
        for c in range(4):
            acc = F(0)
            for k in range(4):
                acc = F(acc + F(a[r, k] * b[k, c]))
            out[r, c] = acc
    return out


def _rotation_axis_angle(axis, angle) -> np.ndarray:
    """3×3 f32 rotation of `UnitQuaternion::from_axis_angle(axis, angle)` (axis already normalised)."""
    half = F(F(angle) * F(0.5))
    s, w = F(np.sin(half)), F(np.cos(half))
    x, y, z = (F(axis[0] * s), F(axis[1] * s), F(axis[2] * s))
    two = F(2)
    return np.array([
        [F(1) - two * (y * y + z * z), two * (x * y - w * z), two * (x * z + w * y)],
        [two * (x * y + w * z), F(1) - two * (x * x + z * z), two * (y * z - w * x)],
        [two * (x * z - w * y), two * (y * z + w * x), F(1) - two * (x * x + y * y)]], dtype=F)


def _homogeneous(rot3: np.ndarray, translation) -> np.ndarray:
    m = np.eye(4, dtype=F)
    m[:3, :3] = rot3
    m[:3, 3] = np.asarray(translation, dtype=F)
    return m


# ---- interpreter (interpreter.rs:100-380) ---------------------------------------------------------------------------
class CardInterpreter:
    def __init__(self, keep_ignored: bool = False):
        self.comments: list = []
        self.ignored: Optional[list] = [] if keep_ignored else None
        self.geometry: dict = {}            # tag → Geometry; iterated in tag order like the reference's BTreeMap
        self.deferred_insertions: list = []
        self.deferred_removals: list = []
        self.symmetry: tuple = ("none",)
        self.ground_plane: Optional[bool] = None

    def finish(self) -> NecFile:
        return NecFile(self.comments, sorted(self.geometry.items(), key=lambda kv: kv[0]), self.ground_plane, self.symmetry,
                       self.ignored or [])

    # -- helpers -------------------------------------------------------------------------------------------------
    def _apply_deferred(self) -> None:       # interpreter.rs:167-174
        for tag in self.deferred_removals:
            self.geometry.pop(tag, None)
        for tag, g in self.deferred_insertions:
            self.geometry[tag] = g
        self.deferred_removals, self.deferred_insertions = [], []

    def _modify(self, tag_increment: int, num_new: int, base: np.ndarray, tag_start: Optional[int], keep_tag: bool) -> None:
        """interpreter.rs:124-165 (GM and GR)."""
        def new_tag(tag: int, i: int) -> int:
            return tag + tag_increment * i if (tag != 0 and not keep_tag) else tag

        for tag in sorted(self.geometry):
            if tag_start is not None and tag < tag_start:
                continue
            g = self.geometry[tag]
            if num_new == 0:
                g.append_transform(base)            # move the structure
                if tag_increment != 0:
                    self.deferred_removals.append(tag)
                    self.deferred_insertions.append((new_tag(tag, 1), g.copy()))
            else:
                dup = g.copy()
                for i in range(num_new):            # cumulative: copy i carries base^(i+1)
                    dup.append_transform(base)
                    self.deferred_insertions.append((new_tag(tag, i + 1), dup.copy()))
        self._apply_deferred()

    # -- cards ---------------------------------------------------------------------------------------------------
    def comment(self, text: str) -> None:
        self.comments.append(text)

    def unknown_card(self, line: str) -> None:
        if self.ignored is not None:
            self.ignored.append(line)

    def wire_arc(self, tag, num_segments, arc_radius, arc_angles, wire_radius) -> None:      # GA, :194-215
        self.geometry[tag] = Geometry("wire_arc", np.eye(4, dtype=F), num_segments=num_segments, arc_radius=arc_radius,
                                      arc_angles=(arc_angles[0], arc_angles[1]), wire_radius=wire_radius)
        self.symmetry = ("none",)

    def end_geometry(self, ground_plane: Optional[bool]) -> None:                             # GE, :218-225
        self.ground_plane = ground_plane
        if ground_plane is not None and self.symmetry[0] == "planar":
            self.symmetry = ("planar", self.symmetry[1] & ~0b100)

    def coordinate_transformation(self, tag_increment, num_new, rotation_deg, translation, tag_start) -> None:   # GM, :228-252
        if num_new > 0 or tag_start is not None:
            self.symmetry = ("none",)
        rad = [F(np.radians(F(a))) for a in rotation_deg]
        rz = _rotation_axis_angle((F(0), F(0), F(1)), rad[2])
        ry = _rotation_axis_angle((F(0), F(1), F(0)), rad[1])
        rx = _rotation_axis_angle((F(1), F(0), F(0)), rad[0])
        rot = (rz @ ry @ rx).astype(F)
        self._modify(tag_increment, num_new, _homogeneous(rot, translation), tag_start, False)

    def cylindrical_structure(self, tag_increment, num_copies) -> None:                        # GR, :255-274
        num_copies = min(num_copies, 1)   # the reference writes `.min(1)` (sic) — at most one copy survives
        if num_copies == 0:
            raise NecError("GR with zero copies (the reference underflows `num_copies - 1` here)")
        angle = F(F(math.tau) / F(num_copies))
        base = _homogeneous(_rotation_axis_angle((F(0), F(0), F(1)), angle), (0, 0, 0))
        self.symmetry = ("rotational",)
        self._modify(tag_increment, num_copies - 1, base, None, True)

    def scale_structure(self, s) -> None:                                                      # GS, :277-308
        m = np.zeros((4, 4), dtype=F)
        m[0, 0] = m[1, 1] = m[2, 2] = F(s)
        m[3, 3] = F(1)
        for g in self.geometry.values():
            g.transform = _matmul4(m, g.transform)
            if g.kind == "wire_arc":
                g.arc_radius = F(g.arc_radius * F(s))
                g.wire_radius = F(g.wire_radius * F(s))
            else:
                g.length = F(g.length * F(s))
                g.segments.scale(F(s))

    def wire(self, tag, num_segments, ends: np.ndarray, segments: WireSegments) -> None:       # GW, :311-335
        delta = (ends[0] - ends[1]).astype(F)
        length = F(np.sqrt(F(F(delta[0] * delta[0]) + F(delta[1] * delta[1]) + F(delta[2] * delta[2]))))
        # `UnitQuaternion::from_axis_angle(normalize(delta), 0.0)` is the identity: the wire's direction is dropped here
        self.geometry[tag] = Geometry("wire", _homogeneous(np.eye(3, dtype=F), ends[0]), length=length,
                                      num_segments=num_segments, segments=segments)
        self.symmetry = ("none",)

    def reflection(self, tag_increment, axes: int) -> None:                                    # GX, :338-371
        # insertions are only *deferred* (the reference never applies them here; the next GM / GR does)
        for bit, axis in ((0b100, 2), (0b010, 1), (0b001, 0)):
            if axes & bit:
                m = np.eye(4, dtype=F)
                m[axis, axis] = F(-1)
                for tag in sorted(self.geometry):
                    g = self.geometry[tag].copy()
                    g.transform = _matmul4(m, g.transform)
                    self.deferred_insertions.append((tag + tag_increment, g))
                tag_increment *= 2
        self.symmetry = ("planar", axes)


# ---- deck parser (parser.rs:84-292) ------------------------------------------------------------------------------------
def _ground_plane_flag(tokens: _Tokens) -> Optional[bool]:
    """card.rs:300-337 read with `.ok().unwrap_or_default()` (parser.rs:142-147): anything unparsable = not present."""
    t = tokens.token()
    try:
        n = int(t) if t is not None else None
    except ValueError:
        n = None
    return {1: True, -1: False}.get(n)


def _reflection_axis(tokens: _Tokens) -> int:
    """card.rs:367-395: exactly three characters '0'/'1' for X, Y, Z."""
    t = tokens._need()
    if len(t) != 3 or any(c not in "01" for c in t):
        raise InvalidParameter(t)
    return (0b001 if t[0] == "1" else 0) | (0b010 if t[1] == "1" else 0) | (0b100 if t[2] == "1" else 0)


def parse_deck(lines: Iterable[str], keep_ignored: bool = False) -> NecFile:
    """`NecFile::from_reader` (interpreter.rs:47-54): reads cards until GE; cards after GE (EX, RP, EN …) are not read
    (parser.rs:91-100) — the reference has no excitation or radiation-pattern handling."""
    it = iter(lines)
    h = CardInterpreter(keep_ignored)
    state: tuple = ("comments",)
    while state[0] != "end":
        section = "comments" if state[0] == "comments" else "geometry"
        try:
            line = next(it).rstrip("\r\n")
        except StopIteration:
            raise UnexpectedEnd(f"deck ended inside the {section} section") from None
        tk = _Tokens(line, section)
        card = tk.card_type()
        if state[0] == "comments":
            if card == "CM":
                h.comment(tk.remainder())
            elif card == "CE":
                if tk.remainder():
                    h.comment(tk.remainder())
                state = ("geometry",)
            else:
                raise UnexpectedCardForSection(f"{card} in the comments section")
        elif state[0] == "geometry":
            if card == "GA":
                h.wire_arc(tk.u32(), tk.u32(), tk.f32(), tk.f32s(2), tk.f32())
            elif card == "GE":
                h.end_geometry(_ground_plane_flag(tk))
                state = ("end",)
            elif card == "GM":
                inc, new = tk.u32(), tk.u32()
                rot, tr = tk.f32s(3), tk.f32s(3)
                start = None
                try:  # optional, read as f32 and rounded; 0 = none (parser.rs:155-160)
                    v = int(np.round(tk.f32()))
                    start = v if 0 < v <= 0xFFFFFFFF else None
                except NecError:
                    start = None
                h.coordinate_transformation(inc, new, rot, tr, start)
            elif card == "GR":
                h.cylindrical_structure(tk.u32(), tk.u32())
            elif card == "GS":
                h.scale_structure(tk.f32())
            elif card == "GW":
                tag, nseg = tk.u32(), tk.u32()
                ends = np.stack([tk.f32s(3), tk.f32s(3)])
                radius = tk.f32()
                if radius == F(0):
                    state = ("gc", tag, nseg, ends)
                else:
                    h.wire(tag, nseg, ends, WireSegments("fixed", radius=radius))
            elif card == "GX":
                h.reflection(tk.u32(), _reflection_axis(tk))
            elif card == "SP":
                raise NotImplementedError("surface patches: `todo!()` in the reference (interpreter.rs:374-380)")
            else:
                h.unknown_card(line)
        elif state[0] == "gc":   # the card after a zero-radius GW: taper parameters read straight after the identifier
            _, tag, nseg, ends = state
            h.wire(tag, nseg, ends, WireSegments("tapered", length_ratio=tk.f32(), first_radius=tk.f32(), last_radius=tk.f32()))
            state = ("geometry",)
    return h.finish()


# ---- deck → colliders (cem-app/src/composer/file_formats/nec.rs:38-82) -------------------------------------------------
@dataclass(frozen=True)
class WireCylinder:
    """One wire segment as the application places it: a parry `Cylinder::new(half_height, radius)` (axis = local Y) with
    identity rotation at `centre` (world units)."""
    tag: int
    segment: int
    centre: tuple          # 3 × float32
    half_height: np.float32
    radius: np.float32


def wire_cylinders(nec: NecFile) -> list:
    """Segment *i* of a wire sits at `transform · (0, i·len, 0, 1)` with an identity rotation (the quaternion is built
    with angle 0), i.e. the reference strings every wire along +Y from its first end regardless of the wire's direction,
    and the first cylinder is *centred* on that end (nec.rs:47-70).  Restated as-is."""
    out = []
    for tag, g in nec.geometry:
        if g.kind != "wire":
            raise NotImplementedError("wire arcs: `todo!()` in the reference (nec.rs:41)")
        for i, dim in enumerate(g.segments.dimensions(g.num_segments, g.length)):
            if dim[0] != "flat":
                raise NotImplementedError("tapered segments: `todo!()` in the reference (nec.rs:73)")
            _, length, radius = dim
            p = np.array([F(0), F(F(i) * length), F(0), F(1)], dtype=F)
            c = [F(0)] * 3
            for r in range(3):
                acc = F(0)
                for k in range(4):
                    acc = F(acc + F(g.transform[r, k] * p[k]))
                c[r] = acc
            out.append(WireCylinder(tag, i, tuple(c), F(F(0.5) * length), F(radius)))
    return out
