"""Seeded synthetic workloads of the shapes BASELINE.json names, encoded to FLS images.

  basic_usage(n)      config 1: the table of /root/reference/examples/basic_usage.sql:11-17 (+ BIGINT, entropy DOUBLE)
  ffor_values(...)    config 2: one FFOR column whose encoder picks exactly width W
  alp_prices / alp_entropy   config 3
  lineitem(n)         configs 4/5: TPC-H lineitem-shaped, 16 columns
                      (schema of /root/reference/duckdb/extension/tpch/dbgen/dbgen.cpp:415-425; DECIMAL -> DOUBLE
                      because the reference's type table has no DECIMAL, /root/reference/src/type_mapping.cpp:58-60)

Generators are numpy-vectorised; encoding goes through the C++ writer (fls_writer.py).  Not on the timed path.
"""
import numpy as np

import fls_writer as W
from fls_writer import StrColumn

RG = W.RG_ROWS


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def _str_from_lens(pool: np.ndarray, starts: np.ndarray, lens: np.ndarray) -> StrColumn:
    """Strings pool[starts[i] : starts[i]+lens[i]] gathered into one Arrow column."""
    off = np.zeros(len(lens) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    total = int(off[-1])
    idx = np.repeat(starts.astype(np.int64) - off[:-1].astype(np.int64), lens) + np.arange(total, dtype=np.int64)
    return StrColumn(pool[idx], off)


def _decimal_strings(prefix: bytes, ids: np.ndarray) -> StrColumn:
    """prefix || str(id) for non-negative ids, vectorised."""
    ids = ids.astype(np.int64)
    nd = np.ones(len(ids), dtype=np.int64)
    p = 10
    while True:
        m = ids >= p
        if not m.any():
            break
        nd += m
        p *= 10
    lens = nd + len(prefix)
    off = np.zeros(len(ids) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    out = np.empty(int(off[-1]), dtype=np.uint8)
    base = off[:-1].astype(np.int64)
    for j, ch in enumerate(prefix):
        out[base + j] = ch
    maxd = int(nd.max()) if len(ids) else 0
    for d in range(maxd):  # digit d counted from the least significant
        m = nd > d
        pos = base[m] + len(prefix) + nd[m] - 1 - d
        out[pos] = (ids[m] // (10 ** d)) % 10 + 48
    return StrColumn(out, off)


# ------------------------------------------------------------------------------------------------ config 1
def basic_usage(n=1_000_000, seed=42):
    r = _rng(seed)
    ids = np.arange(1, n + 1, dtype=np.int32)
    big = ids.astype(np.int64) * 1_000_003 + 1_234_567_890_123
    score = np.round(r.uniform(0, 1000, n), 2)
    rnd = r.uniform(0, 1, n) * 1000.0
    user = _decimal_strings(b"user_", ids)
    active = (r.integers(0, 2, n) == 1)
    date = (19358 + (ids % 365)).astype(np.int32)  # 2023-01-01 + id % 365
    cols = [ids, big, score, rnd, user, active, date]
    names = ["id", "big", "score", "rnd", "username", "is_active", "join_date"]
    return cols, names


def basic_usage_types():
    return [W.INT32, W.INT64, W.DOUBLE, W.DOUBLE, W.STR, W.BOOLEAN, W.DATE]


# ------------------------------------------------------------------------------------------------ config 2
def ffor_values(bits, width, n, seed=None):
    """n values of a `bits`-wide integer type whose every 1024-vector needs exactly `width` FFOR bits."""
    seed = 0xF15000 + 100 * (bits // 32) + width if seed is None else seed
    r = _rng(seed)
    udt = {8: np.uint8, 16: np.uint16, 32: np.uint32, 64: np.uint64}[bits]
    sdt = {8: np.int8, 16: np.int16, 32: np.int32, 64: np.int64}[bits]
    if width == 0:
        v = np.zeros(n, dtype=udt)
    elif width == 64:
        v = r.integers(0, 2 ** 64, n, dtype=np.uint64, endpoint=False)
    else:
        v = r.integers(0, 2 ** width, n, dtype=np.uint64, endpoint=False).astype(udt)
    nv = (n + 1023) // 1024
    if width:
        v[np.arange(nv) * 1024] = 0  # min and max present in every vector -> exactly `width` bits
        last = np.minimum(np.arange(nv) * 1024 + 1, n - 1)
        v[last] = udt(2 ** width - 1)
    if bits == 32:
        base = 1_000_003 if width < 32 else 0
    elif bits == 64:
        base = -(2 ** 40) if width < 64 else 0
    else:
        base = 3 if width < bits else 0
    return (v + udt(base % (2 ** bits))).view(sdt)


# ------------------------------------------------------------------------------------------------ config 3
def alp_prices(n, seed=0xA1B, exc_rate=0.001):
    r = _rng(seed)
    v = np.round(r.uniform(1, 100000, n) * 100) / 100
    if exc_rate:
        k = max(1, int(n * exc_rate))
        pos = r.choice(n, k, replace=False)
        specials = np.array([np.nan, np.inf, -np.inf, -0.0, 1e300, 0.1 + 0.2])
        v[pos] = specials[r.integers(0, len(specials), k)]
    return v


def alp_entropy(n, seed=0xA1C):
    r = _rng(seed)
    mant = r.integers(0, 2 ** 52, n, dtype=np.uint64)
    # U(0.5,1) x small exponent spread, 52 random mantissa bits
    expo = (np.uint64(1022) - r.integers(0, 4, n, dtype=np.uint64)) << np.uint64(52)
    return (expo | mant).view(np.float64)


# ------------------------------------------------------------------------------------------------ configs 4/5
_WORDS = (b"furiously quickly carefully slyly blithely final special regular express bold ironic pending even unusual "
          b"silent ruthless packages requests deposits accounts instructions theodolites foxes pinto beans dependencies "
          b"excuses platelets asymptotes courts dolphins multipliers sauternes warthogs frets dinos attainments somas "
          b"Tiresias patterns forges braids hockey players frays warhorses dugouts notornis epitaphs pearls tithes "
          b"waters orbits gifts sheaves depths sentiments decoys realms pains grouches escapades sleep wake are cajole "
          b"haggle nag use boost affix detect integrate maintain nod was lose sublate solve thrash promise engage hinder "
          b"print x-ray breach eat grow impress mold poach serve run dazzle snooze doze unwind kindle play hang believe "
          b"doubt about above according to across after against along alongside of among around at atop before behind "
          b"beneath beside besides between beyond by despite during except for from in place of inside instead of into "
          b"near of on outside over past since through throughout to toward under until up upon without with within ")
LINEITEM_NAMES = ["l_orderkey", "l_partkey", "l_suppkey", "l_linenumber", "l_quantity", "l_extendedprice", "l_discount",
                  "l_tax", "l_returnflag", "l_linestatus", "l_shipdate", "l_commitdate", "l_receiptdate",
                  "l_shipinstruct", "l_shipmode", "l_comment"]
LINEITEM_TYPES = [W.INT64, W.INT64, W.INT64, W.INT64, W.DOUBLE, W.DOUBLE, W.DOUBLE, W.DOUBLE, W.STR, W.STR, W.DATE,
                  W.DATE, W.DATE, W.STR, W.STR, W.STR]
_INSTRUCT = [b"DELIVER IN PERSON", b"COLLECT COD", b"NONE", b"TAKE BACK RETURN"]
_MODES = [b"REG AIR", b"AIR", b"RAIL", b"SHIP", b"TRUCK", b"MAIL", b"FOB"]


def _pick(choices, idx):
    lens = np.array([len(c) for c in choices], dtype=np.int64)
    starts = np.zeros(len(choices), dtype=np.int64)
    starts[1:] = np.cumsum(lens)[:-1]
    pool = np.frombuffer(b"".join(choices), dtype=np.uint8)
    return _str_from_lens(pool, starts[idx], lens[idx])


def lineitem(n, seed=7, first_order=0, text_pool=None):
    """n rows of lineitem-shaped data (orders of 1..7 lines, keys sparse-ascending as dbgen makes them)."""
    r = _rng(seed)
    # orders: enough orders to cover n lines
    n_ord = n // 4 + 16
    cnt = r.integers(1, 8, n_ord)
    while cnt.sum() < n:
        cnt = np.concatenate([cnt, r.integers(1, 8, n_ord)])
    ordn = np.repeat(np.arange(len(cnt), dtype=np.int64), cnt)[:n] + first_order
    start = np.zeros(len(cnt) + 1, dtype=np.int64)
    np.cumsum(cnt, out=start[1:])
    linenumber = (np.arange(n, dtype=np.int64) - np.repeat(start[:-1], cnt)[:n]) + 1
    orderkey = (ordn // 8) * 32 + ordn % 8 + 1  # dbgen's sparse key: 8 of every 32 values are used
    partkey = r.integers(1, 2_000_001, n, dtype=np.int64)
    suppkey = r.integers(1, 100_001, n, dtype=np.int64)
    qty_i = r.integers(1, 51, n, dtype=np.int64)
    retail_c = 90000 + (partkey // 10) % 20001 + 100 * (partkey % 1000)  # cents
    quantity = qty_i.astype(np.float64)
    extendedprice = (qty_i * retail_c).astype(np.float64) / 100.0
    discount = r.integers(0, 11, n).astype(np.float64) / 100.0
    tax = r.integers(0, 9, n).astype(np.float64) / 100.0
    odate_per_order = r.integers(8035, 10441, len(cnt))  # 1992-01-01 .. 1998-08-02
    odate = np.repeat(odate_per_order, cnt)[:n]
    shipdate = (odate + r.integers(1, 122, n)).astype(np.int32)
    commitdate = (odate + r.integers(30, 91, n)).astype(np.int32)
    receiptdate = (shipdate + r.integers(1, 31, n)).astype(np.int32)
    today = 9298  # 1995-06-17
    rf = np.where(receiptdate <= today, np.where(r.integers(0, 2, n) == 0, 0, 2), 1)  # A/R vs N
    returnflag = _pick([b"A", b"N", b"R"], rf)
    linestatus = _pick([b"O", b"F"], (shipdate <= today).astype(np.int64))
    shipinstruct = _pick(_INSTRUCT, r.integers(0, 4, n))
    shipmode = _pick(_MODES, r.integers(0, 7, n))
    if text_pool is None:
        words = _WORDS.split()
        wi = _rng(12345).integers(0, len(words), 40000)
        text_pool = np.frombuffer(b" ".join(words[i] for i in wi), dtype=np.uint8)
    clen = r.integers(10, 44, n)
    cstart = r.integers(0, len(text_pool) - 64, n)
    comment = _str_from_lens(text_pool, cstart, clen)
    return [orderkey, partkey, suppkey, linenumber, quantity, extendedprice, discount, tax, returnflag, linestatus,
            shipdate, commitdate, receiptdate, shipinstruct, shipmode, comment]


# ------------------------------------------------------------------------------------------------ encoding
def encode(cols, names=None, types=None, force=None, min_bw=None, path=None):
    return W.write_table(cols, names=names, types=types, force=force, min_bw=min_bw, path=path)


def lineitem_image(n_rowgroups, seed=7, distinct=None, tail_rows=0):
    """FLS image of `n_rowgroups` full row groups (+ an optional partial one) of lineitem-shaped data.

    distinct: generate and encode only this many distinct row groups and repeat them (each repeat is its own
    copy of the bytes in the file, so nothing is shared in cache); None = all distinct.
    """
    distinct = n_rowgroups if distinct is None else min(distinct, n_rowgroups)
    cols = lineitem(distinct * RG + (tail_rows if distinct == n_rowgroups else 0), seed)
    img = encode(cols, LINEITEM_NAMES, LINEITEM_TYPES)
    if distinct < n_rowgroups:
        img = W.tile_rowgroups(img, n_rowgroups)
    return img
