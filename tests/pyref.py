"""Independent pure-Python restatements used to pin the C oracle (tests only).

Written from the recurrence in natural (i, j) matrix coordinates, NOT from the
rolling-array form of oracle.c, so a transcription slip in either shows up as a
disagreement.  Follows SURVEY.md Appendix A.1 / A.2.
"""


def extend_py(query, target, mat, o_del, e_del, o_ins, e_ins, w, end_bonus, zdrop, h0):
    qlen, tlen = len(query), len(target)
    oe_del, oe_ins = o_del + e_del, o_ins + e_ins
    maxsc = max(max(mat), 0)
    max_ins = max(int((qlen * maxsc + end_bonus - o_ins) / e_ins + 1.0), 1)
    max_del = max(int((qlen * maxsc + end_bonus - o_del) / e_del + 1.0), 1)
    w = min(w, max_ins, max_del)
    # "memory" of the rolling arrays, in natural coordinates:
    #   Hprev[j] = value a later row finds as its diagonal predecessor of column j (H of column j-1)
    #   Ecur[j]  = E entering the next row at column j
    Hprev = [0] * (qlen + 2)
    Ecur = [0] * (qlen + 2)
    Hprev[0] = h0
    if qlen >= 1:
        Hprev[1] = h0 - oe_ins if h0 > oe_ins else 0
    j = 2
    while j <= qlen and Hprev[j - 1] > e_ins:
        Hprev[j] = Hprev[j - 1] - e_ins
        j += 1
    best, best_i, best_j, best_ie, gscore, max_off = h0, -1, -1, -1, -1, 0
    beg, end, cells = 0, qlen, 0
    for i in range(tlen):
        beg = max(beg, i - w)
        end = min(end, i + w + 1, qlen)
        left = max(h0 - (o_del + e_del * (i + 1)), 0) if beg == 0 else 0   # H(i, beg-1)
        f, m, mj = 0, 0, -1
        cells += max(end - beg, 0)
        row = {}
        for j in range(beg, end):
            diag, e = Hprev[j], Ecur[j]
            M = diag + mat[target[i] * 5 + query[j]] if diag else 0
            h = max(M, e, f)
            row[j] = h
            if h >= m:
                mj = j
            m = max(m, h)
            Ecur[j] = max(e - e_del, M - oe_del, 0)
            f = max(f - e_ins, M - oe_ins, 0)
        # commit the row: column j's H becomes the diagonal predecessor of column j+1
        prev = left
        for j in range(beg, end):
            Hprev[j] = prev
            prev = row[j]
        Hprev[end] = prev
        Ecur[end] = 0
        last = prev
        after = end if beg < end else beg
        if after == qlen:
            if not gscore > last:
                best_ie = i
            gscore = max(gscore, last)
        if m == 0:
            break
        if m > best:
            best, best_i, best_j = m, i, mj
            max_off = max(max_off, abs(mj - i))
        elif zdrop > 0:
            di, dj = i - best_i, mj - best_j
            if di > dj:
                if best - m - (di - dj) * e_del > zdrop:
                    break
            elif best - m - (dj - di) * e_ins > zdrop:
                break
        j = beg
        while j < end and Hprev[j] == 0 and Ecur[j] == 0:
            j += 1
        beg = j
        j = end
        while j >= beg and Hprev[j] == 0 and Ecur[j] == 0:
            j -= 1
        end = min(j + 2, qlen)
    return dict(score=best, qle=best_j + 1, tle=best_i + 1, gtle=best_ie + 1, gscore=gscore,
                max_off=max_off, cells=cells)


def ksw_pass_py(query, target, mat, o_del, e_del, o_ins, e_ins, xtra, is8):
    """Full-matrix local SW + ksw_u8's bookkeeping (A.2)."""
    XSTOP, XSUBO = 0x20000, 0x40000
    qlen, tlen = len(query), len(target)
    oe_del, oe_ins = o_del + e_del, o_ins + e_ins
    shift, mx = abs(min(mat)), max(mat)
    minsc = (xtra & 0xffff) if xtra & XSUBO else 0x10000
    endsc = (xtra & 0xffff) if xtra & XSTOP else 0x10000
    H = [[0] * (qlen + 1) for _ in range(tlen + 1)]       # H[i+1][j+1] = H(i, j)
    E = [0] * (qlen + 1)
    gmax, te, blist, rows, hmax_row = 0, -1, [], 0, None
    for i in range(tlen):
        f, imax = 0, 0
        rows += 1
        for j in range(qlen):
            h = max(H[i][j] + mat[target[i] * 5 + query[j]], E[j], f, 0)
            H[i + 1][j + 1] = h
            imax = max(imax, h)
            E[j] = max(E[j] - e_del, h - oe_del, 0)
            f = max(f - e_ins, h - oe_ins, 0)
        if imax >= minsc:
            if not blist or blist[-1][1] + 1 != i:
                blist.append([imax, i])
            elif blist[-1][0] < imax:
                blist[-1] = [imax, i]
        if imax > gmax:
            gmax, te, hmax_row = imax, i, H[i + 1][1:]
            if (is8 and gmax + shift >= 255) or gmax >= endsc:
                break
    ovf = bool(is8 and gmax + shift >= 255)
    r = dict(score=255 if ovf else gmax, te=te, qe=-1, score2=-1, te2=-1, cells=rows * qlen)
    if not ovf:
        if qlen > 0:
            row = hmax_row if hmax_row is not None else [0] * qlen
            r["qe"] = row.index(max(row))
        d = (r["score"] + mx - 1) // mx
        for sc, row_i in blist:
            if (row_i < te - d or row_i > te + d) and sc > r["score2"]:
                r["score2"], r["te2"] = sc, row_i
    return r


def global_full(query, target, mat, o_del, e_del, o_ins, e_ins, w):
    """Independent full-matrix transcription of the recurrences ksw_global2 documents (bwa ksw.c): three (tlen+1) x (qlen+1)
    tables, no rolling row, no backtrack bytes.  Cell (I, J) (1-based) exists iff |I - J| <= w; gaps open from M only.
    Returns H[tlen][qlen]."""
    NEG = -0x40000000
    tl, ql = len(target), len(query)
    H = [[NEG] * (ql + 1) for _ in range(tl + 1)]
    E = [[NEG] * (ql + 1) for _ in range(tl + 2)]     # E[I][J]: gap in the query direction arriving at row I (deletion)
    H[0][0] = 0
    for J in range(1, ql + 1):
        if J <= w:
            H[0][J] = -(o_ins + e_ins * J)
    for I in range(1, tl + 1):
        if I <= w:
            H[I][0] = -(o_del + e_del * I)
    for I in range(1, tl + 1):
        F = NEG
        for J in range(1, ql + 1):
            if abs(I - J) > w:
                # ksw_global2 evaluates j in [i - w, i + w + 1) with i = I - 1, j = J - 1
                continue
            diag = H[I - 1][J - 1]
            M = diag + mat[target[I - 1] * 5 + query[J - 1]]
            e = E[I][J]
            if J - 1 == max(I - 1 - w, 0):          # first cell of the row: F restarts
                F = NEG
            h = max(M, e, F)
            H[I][J] = h
            E[I + 1][J] = max(M - o_del, e) - e_del
            F = max(M - o_ins, F) - e_ins
    return H[tl][ql]


def cigar_score(ops, query, target, mat, o_del, e_del, o_ins, e_ins):
    """Score of an alignment given as bwa ops (len << 4 | op; 0 = M, 1 = I, 2 = D) under affine gaps."""
    i = j = 0
    s = 0
    for o in ops:
        ln, op = int(o) >> 4, int(o) & 15
        if op == 0:
            for _ in range(ln):
                s += mat[target[i] * 5 + query[j]]; i += 1; j += 1
        elif op == 1:
            s -= o_ins + e_ins * ln; j += ln
        else:
            s -= o_del + e_del * ln; i += ln
    assert i == len(target) and j == len(query)
    return s
