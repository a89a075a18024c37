"""Orientation-tagged positions. Mirrors src/pos.rs:6-87 of the reference:
``Pos = u64``, offset in the high bits, orientation (1 = reverse) in the least significant bit."""


def make_pos(offset: int, is_reverse: bool) -> int:  # src/pos.rs:10-12
    return (offset << 1) | (1 if is_reverse else 0)


def is_rev(pos: int) -> bool:  # :16-18
    return (pos & 1) == 1


def offset(pos: int) -> int:  # :22-24
    return pos >> 1


def incr_pos(pos: int) -> int:  # :28-41 (the reverse strand moves towards smaller offsets, clamped at 0)
    if is_rev(pos):
        off = offset(pos)
        return make_pos(off - 1, True) if off > 0 else pos
    return make_pos(offset(pos) + 1, False)


def decr_pos(pos: int) -> int:  # :45-58
    if is_rev(pos):
        return make_pos(offset(pos) + 1, True)
    off = offset(pos)
    return make_pos(off - 1, False) if off > 0 else pos


def flip_orientation(pos: int) -> int:  # :62-64
    return pos ^ 1


def orientation_char(pos: int) -> str:  # :68-74
    return "-" if is_rev(pos) else "+"


_RC = {ord("A"): ord("T"), ord("a"): ord("T"), ord("T"): ord("A"), ord("t"): ord("A"),
       ord("C"): ord("G"), ord("c"): ord("G"), ord("G"): ord("C"), ord("g"): ord("C"),
       ord("N"): ord("N"), ord("n"): ord("N")}


def rc_base(base: int) -> int:  # :78-87
    return _RC.get(base, base)
