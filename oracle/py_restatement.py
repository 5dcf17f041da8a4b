"""Independent pure-Python restatement of the reference recurrence — TEST INFRASTRUCTURE.

Second opinion for oracle/*.cpp (differential tests on small cases); written separately
from the C++ text, straight from the Rust lines cited.  Never imported by the product.

Citations relative to /root/reference/.
"""
import math

TOP, LEFT, DIAGONAL, BEGINNING = 0, 1, 2, 3
BLANK = 98
EPS = 2.220446049250313e-16  # f64::EPSILON


def _fmax(a, b):
    # Rust f64::max: if one operand is NaN return the other
    if a != a:
        return b
    if b != b:
        return a
    return a if a > b else b


def get_direction(top, left, diag, with_beginning):
    """aligner-core/src/enums.rs:18-28 and :30-46"""
    mx = _fmax(_fmax(top, left), diag)
    if with_beginning and mx == 0.0:
        return mx, BEGINNING
    if abs(mx - top) < EPS:
        return mx, TOP
    if abs(mx - left) < EPS:
        return mx, LEFT
    return mx, DIAGONAL


def align(mode, q, t, dele, ext, matrix):
    """mode 'global' = simple/mod.rs:42-144; 'local' = simple/mod.rs:168-264.

    matrix: list of rows, indexed matrix[t_sym][q_sym].  Returns dict with H, D (lists of
    rows), q_aln, t_aln, coords ((a,b),(c,d)), f.  Raises ValueError where Rust panics.
    """
    n, m = len(q), len(t)
    if n == 0 or m == 0:
        raise ValueError("empty")
    H = [[0.0] * (n + 1) for _ in range(m + 1)]
    D = [[BEGINNING] * (n + 1) for _ in range(m + 1)]
    glob = mode == "global"
    if glob:
        for x in range(1, n + 1):
            H[0][x] = -float(x) * dele
            D[0][x] = LEFT
        for y in range(1, m + 1):
            H[y][0] = -float(y) * dele
            D[y][0] = TOP
        H[0][n] = -(float(n) + 1.0) * dele
        H[m][0] = -(float(m) + 1.0) * dele
    penalty = dele
    for x in range(1, n + 1):
        for y in range(1, m + 1):
            v, d = get_direction(
                H[y - 1][x] - penalty,
                H[y][x - 1] - penalty,
                H[y - 1][x - 1] + matrix[t[y - 1]][q[x - 1]],
                not glob,
            )
            penalty = ext if d != BEGINNING else dele
            H[y][x] = v
            D[y][x] = d
    if glob:
        sy, sx = m, n
        f = 0.0
    else:
        best, sy, sx = H[0][0], 0, 0
        for y in range(m + 1):
            row = H[y]
            for x in range(n + 1):
                if row[x] > best:
                    best, sy, sx = row[x], y, x
        if sy == 0 or sx == 0:
            raise ValueError("no positive cell")
        f = best
    cy, cx = sy, sx
    qa, ta = [q[cx - 1]], [t[cy - 1]]
    while True:
        d = D[cy][cx]
        if d == BEGINNING:
            break
        if d == TOP:
            qa.append(BLANK)
            ta.append(t[cy - 1])
            cy -= 1
        elif d == LEFT:
            qa.append(q[cx - 1])
            ta.append(BLANK)
            cx -= 1
        else:
            qa.append(q[cx - 1])
            ta.append(t[cy - 1])
            cx -= 1
            cy -= 1
    qa.reverse()
    ta.reverse()
    coords = ((1, n), (1, m)) if glob else ((cx + 1, sx + 1), (cy + 1, sy + 1))
    return dict(H=H, D=D, q_aln=qa, t_aln=ta, coords=coords, f=f)


def frequency_matrix(qa, ta, V):
    """aligner-core/src/alignment.rs:13-23"""
    F = [[0.0] * V for _ in range(V)]
    for x, y in zip(qa, ta):
        if x != BLANK and y != BLANK:
            F[y][x] += 1.0
    return F


def _nd_sum(xs):
    """ndarray 0.15.4 unrolled_fold order (third-party, restated from memory)."""
    p = [0.0] * 8
    i = 0
    while len(xs) - i >= 8:
        for k in range(8):
            p[k] = p[k] + xs[i + k]
        i += 8
    acc = 0.0
    acc = acc + (p[0] + p[4])
    acc = acc + (p[1] + p[5])
    acc = acc + (p[2] + p[6])
    acc = acc + (p[3] + p[7])
    for x in xs[i:]:
        acc = acc + x
    return acc


def _roots_quadratic(a2, a1, a0):
    """roots 0.0.7 find_roots_quadratic (third-party, restated from memory)."""
    disc = a1 * a1 - 4.0 * a2 * a0
    if disc < 0.0:
        return []
    a2x2 = 2.0 * a2
    if disc == 0.0:
        return [-a1 / a2x2]
    sq = math.sqrt(disc)
    if a1 < 0.0:
        same, diff = -a1 + sq, -a1 - sq
    else:
        same, diff = -a1 - sq, -a1 + sq
    if abs(same) > abs(a2x2):
        a0x2 = 2.0 * a0
        if abs(diff) > abs(a2x2):
            x1, x2 = a0x2 / same, a0x2 / diff
        else:
            x1, x2 = a0x2 / same, same / a2x2
    else:
        x1, x2 = diff / a2x2, same / a2x2
    return [x1, x2] if x1 < x2 else [x2, x1]


def transform_matrix(M, k_d, r_squared, freq):
    """aligner-helpers/src/matrices/mod.rs:19-68; M list of rows. Raises ValueError on Roots::No."""
    R, C = len(M), len(M[0])
    flat = [v for row in M for v in row]
    fcol = 1.0 / float(C)
    p = [freq[y] * fcol for y in range(R) for _ in range(C)]
    p_squared = _nd_sum([a * a for a in p])
    k_0 = _nd_sum([a * b for a, b in zip(p, flat)])
    a = (k_d - k_0) / p_squared
    b = k_d / p_squared
    diff = a - b
    e = [mv + pv * diff for mv, pv in zip(flat, p)]
    den = _nd_sum([v * v for v in e])
    a_c = (2.0 * b * _nd_sum([pv * ev for pv, ev in zip(p, e)])) / den
    b_c = (b * b * p_squared - r_squared) / den
    roots = _roots_quadratic(1.0, a_c, b_c)
    if not roots:
        raise ValueError("WrongMatrixSpecified")

    def build(r):
        return [pv * b + r * (mv + pv * (a - b)) for pv, mv in zip(p, flat)]

    if len(roots) == 1:
        out = build(roots[0])
    elif roots[0] > 0.0 and roots[1] < 0.0:
        out = build(roots[0])
    elif roots[0] < 0.0 and roots[1] > 0.0:
        out = build(roots[1])
    else:
        m1, m2 = build(roots[0]), build(roots[1])

        def l2(u, v):
            s = 0.0
            for x, y in zip(u, v):
                d = x - y
                s += d * d
            return math.sqrt(s)

        out = m1 if l2(flat, m1) < l2(flat, m2) else m2
    return [out[y * C:(y + 1) * C] for y in range(R)]


def heuristic(q, t, dele, ext, matrix, kd, r_squared, freq):
    """aligner-core/src/heuristic/mod.rs:36-78. Returns (result dict, final matrix, passes)."""
    V = len(matrix)
    if abs(r_squared - 0.0) < EPS:
        r_squared = float(len(matrix) * len(matrix[0]))
    T = transform_matrix(matrix, kd, r_squared, freq)
    max_f = 0.0
    passes = 0
    while True:
        res = align("local", q, t, dele, ext, T)
        passes += 1
        if res["f"] > max_f:
            max_f = res["f"]
            T = transform_matrix(frequency_matrix(res["q_aln"], res["t_aln"], V), kd, r_squared, freq)
        else:
            return res, T, passes


def pwm_align(query, pwm, dele, ext):
    """aligner-core/src/pwm/mod.rs:29-126.  query: window symbols (rows); pwm: 4 rows of L scores.
    Returns dict with H, D, numbered, q_aln, coords ((a,b),(c,d)), f."""
    if len(pwm) != 4:
        raise ValueError("MatrixShapeError")           # :40-42
    L = len(pwm[0])
    numbered_sequence = list(range(1, L + 1))          # :44
    m = len(query)
    H = [[0.0] * (L + 1) for _ in range(m + 1)]        # :48
    D = [[BEGINNING] * (L + 1) for _ in range(m + 1)]  # :49-50
    penalty = dele                                     # :52
    for elem_1 in numbered_sequence:                   # :54
        for y, elem_2 in enumerate(query):             # :55
            y_real = y + 1
            v, d = get_direction(H[y_real - 1][elem_1] - penalty, H[y_real][elem_1 - 1] - penalty,
                                 H[y_real - 1][elem_1 - 1] + pwm[elem_2][elem_1 - 1], True)  # :59-63
            penalty = ext if d != BEGINNING else dele  # :65-69
            H[y_real][elem_1] = v
            D[y_real][elem_1] = d
    best, my, mx = H[0][0], 0, 0                       # :76 argmax: first maximum, row-major
    for y in range(m + 1):
        for x in range(L + 1):
            if H[y][x] > best:
                best, my, mx = H[y][x], y, x
    numbered, qa = [], []
    cx, cy = mx, my
    while True:                                        # :81-104
        d = D[cy][cx]
        if d == BEGINNING:
            break
        if d == TOP:
            numbered.append(0)
            qa.append(query[cy - 1])
            cy -= 1
        elif d == LEFT:
            numbered.append(numbered_sequence[cx - 1])
            qa.append(BLANK)
            cx -= 1
        else:
            numbered.append(numbered_sequence[cx - 1])
            qa.append(query[cy - 1])
            cx -= 1
            cy -= 1
    numbered.reverse()
    qa.reverse()
    return dict(H=H, D=D, numbered=numbered, q_aln=qa, f=best,
                coords=((cx + 1, mx + 1), (cy + 1, my + 1)))  # :118-121


def pwm_frequency_matrix(numbered, qa, L):
    """aligner-core/src/alignment.rs:55-65"""
    F = [[0.0] * L for _ in range(4)]
    for x, y in zip(numbered, qa):
        if x != 0 and y != BLANK:
            F[y][x - 1] += 1.0
    return F
