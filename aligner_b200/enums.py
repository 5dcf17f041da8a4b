"""Symbol types and error kinds of the reference, as numeric codes.

Mirrors aligner-core/src/enums.rs: ``Direction`` (:9-15), ``Protein`` (:55-84),
``DNA`` (:138-147) and the ``BioData`` conversions (:181-199, :202-303, :404-467), and
``aligner_core::Error`` (aligner-core/src/lib.rs:47-57).  Sequences are ``numpy.uint8`` arrays
of codes — exactly what crosses the C ABI.
"""
import numpy as np


class Direction:
    Top, Left, Diagonal, Beginning = 0, 1, 2, 3
    NAMES = ("Top", "Left", "Diagonal", "Beginning")


class Error(Exception):
    """aligner_core::Error; ``kind`` is the variant name (lib.rs:47-57)."""

    KINDS = ("ProteinNotFound", "CharIsNotMatchable", "UnnecessaryArgument", "MissingArgument",
             "ResultIsEmpty", "CalculationError", "ValidationError", "MatrixShapeError")

    def __init__(self, kind, detail=""):
        self.kind = kind
        super().__init__(kind + ((": " + detail) if detail else ""))


class _BioData:
    LETTERS = ""
    BLANK, POS, ANY = 98, 99, 100
    SKIP_UNKNOWN_IN_U8 = False  # DNA::from_u8_vec skips unknown bytes (enums.rs:454-467)

    @classmethod
    def volume(cls):
        return len(cls.LETTERS)

    @classmethod
    def blank(cls):
        return cls.BLANK

    @classmethod
    def pos(cls):
        return cls.POS

    @classmethod
    def match_with_char(cls, ch):
        i = cls.LETTERS.find(ch) if len(ch) == 1 else -1
        if i >= 0:
            return i
        if ch == "_":
            return cls.BLANK
        if ch == "+":
            return cls.POS
        raise Error("CharIsNotMatchable", repr(ch))

    @classmethod
    def convert_to_char(cls, code):
        code = int(code)
        if 0 <= code < len(cls.LETTERS):
            return cls.LETTERS[code]
        if code == cls.BLANK:
            return "_"
        if code == cls.POS:
            return "+"
        if code == cls.ANY:
            return "*"
        raise Error("ProteinNotFound", str(code))

    @classmethod
    def _table(cls):
        tab = cls.__dict__.get("_TAB")
        if tab is None:
            tab = np.full(256, 255, dtype=np.uint8)
            for i, ch in enumerate(cls.LETTERS):
                tab[ord(ch)] = i
            tab[ord("_")] = cls.BLANK
            tab[ord("+")] = cls.POS
            cls._TAB = tab
        return tab

    @classmethod
    def str_to_vec(cls, s):
        """BioData::str_to_vec: any unknown char is an error (enums.rs:266-277, :428-439)."""
        raw = np.frombuffer(s.encode("latin-1", "replace"), dtype=np.uint8)
        out = cls._table()[raw]
        if (out == 255).any():
            raise Error("CharIsNotMatchable", repr(s[int(np.argmax(out == 255))]))
        return out

    @classmethod
    def from_u8_vec(cls, data):
        raw = np.frombuffer(bytes(data), dtype=np.uint8)
        out = cls._table()[raw]
        bad = out == 255
        if bad.any():
            if cls.SKIP_UNKNOWN_IN_U8:
                return out[~bad]
            raise Error("CharIsNotMatchable", chr(int(raw[int(np.argmax(bad))])))
        return out

    @classmethod
    def vec_to_str(cls, codes):
        return "".join(cls.convert_to_char(c) for c in codes)


class Protein(_BioData):
    LETTERS = "ARNDCQEGHILKMFPSTWYVBJZX"  # codes 0..23 (enums.rs:55-84)


class DNA(_BioData):
    LETTERS = "ATCG"  # A=0 T=1 C=2 G=3 (enums.rs:138-147) — not ACGT order
    SKIP_UNKNOWN_IN_U8 = True
