"""Minimal reader for the SVG path mini-language (the `d` attribute), enough for SketchLoader.preprocess
(reference utils.py:124-217), which uses the third-party `svg.path` package only to turn KanjiVG path strings
into line and cubic-Bezier segments with absolute end points.  This is an independent, table-driven
implementation (the vendored svg.path does not import on Python 3.12 and is not copied here); it follows the SVG
1.1 path grammar the way that package does where the two could differ:

  * the first moveto is absolute even when written `m`; further coordinate pairs after a moveto are linetos;
  * `Z` emits a line back to the sub-path's start (even a zero-length one) and forbids implicit repeats;
  * `S` / `T` reflect the previous control point only when the previous COMMAND LETTER was C/S resp. Q/T;
  * numbers may run together (`1.5-2`, `.5.5`), commas and whitespace are interchangeable.

segments(d) returns tuples of complex numbers:
  ("L", start, end)   ("C", start, control1, control2, end)   ("Q", start, control, end)   ("A", start, end)
"""
import re

_NARGS = {"M": 2, "Z": 0, "L": 2, "H": 1, "V": 1, "C": 6, "S": 4, "Q": 4, "T": 2, "A": 7}
_TOKEN = re.compile(r"([MmZzLlHhVvCcSsQqTtAa])|([-+]?[0-9]*\.?[0-9]+(?:[eE][-+]?[0-9]+)?)")


def _tokens(d):
    for m in _TOKEN.finditer(d):
        yield (m.group(1), None) if m.group(1) else (None, float(m.group(2)))


def segments(d):
    out = []
    cur = 0j           # current point
    sub_start = None   # start of the current sub-path (target of Z)
    letter = None      # command in force, upper case
    relative = False
    prev_letter = None  # the command LETTER read before the current one (S / T look at it)
    args = []
    first_move = True

    def run(cmd, a):
        nonlocal cur, sub_start, first_move
        off = cur if relative else 0j
        if cmd == "M":
            p = complex(a[0], a[1])
            cur = p if (first_move or not relative) else cur + p
            first_move = False
            sub_start = cur
        elif cmd == "L":
            p = complex(a[0], a[1]) + off
            out.append(("L", cur, p))
            cur = p
        elif cmd == "H":
            p = complex(a[0] + (cur.real if relative else 0.0), cur.imag)
            out.append(("L", cur, p))
            cur = p
        elif cmd == "V":
            p = complex(cur.real, a[0] + (cur.imag if relative else 0.0))
            out.append(("L", cur, p))
            cur = p
        elif cmd == "C":
            c1, c2, p = complex(a[0], a[1]) + off, complex(a[2], a[3]) + off, complex(a[4], a[5]) + off
            out.append(("C", cur, c1, c2, p))
            cur = p
        elif cmd == "S":
            c1 = 2 * cur - out[-1][3] if (prev_letter in ("C", "S") and out and out[-1][0] == "C") else cur
            c2, p = complex(a[0], a[1]) + off, complex(a[2], a[3]) + off
            out.append(("C", cur, c1, c2, p))
            cur = p
        elif cmd == "Q":
            c, p = complex(a[0], a[1]) + off, complex(a[2], a[3]) + off
            out.append(("Q", cur, c, p))
            cur = p
        elif cmd == "T":
            c = 2 * cur - out[-1][2] if (prev_letter in ("Q", "T") and out and out[-1][0] == "Q") else cur
            p = complex(a[0], a[1]) + off
            out.append(("Q", cur, c, p))
            cur = p
        elif cmd == "A":
            p = complex(a[5], a[6]) + off
            out.append(("A", cur, p))
            cur = p

    for sym, num in _tokens(d):
        if sym is not None:
            if args:
                raise ValueError("path data ends a command with %d stray numbers: %r" % (len(args), d))
            prev_letter, letter, relative = letter, sym.upper(), sym.islower()
            if letter == "Z":
                out.append(("L", cur, sub_start if sub_start is not None else cur))
                cur = sub_start if sub_start is not None else cur
                sub_start = None
                letter = None  # no implicit repeats after a close
            continue
        if letter is None:
            raise ValueError("number without a command in path data %r" % d)
        args.append(num)
        if len(args) == _NARGS[letter]:
            run(letter, args)
            args = []
            if letter == "M":
                letter = "L"  # further pairs after a moveto are linetos
    if args:
        raise ValueError("path data ends in the middle of a command: %r" % d)
    return out
