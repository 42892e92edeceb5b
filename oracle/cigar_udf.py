"""ORACLE (test infrastructure, not product): the reference's CIGAR scalar UDFs, restated in Python on the CIGAR TEXT the way
src/kmer_udf.c computes them -- parse_cigar_metrics (:197-269), cigar_has_operator_text (:271-293), the dispatch of
cigar_metric_scalar / cigar_has_op_scalar (:695-790).  None = the SQL NULL the UDF returns.

Pinned by tests/test_oracle.py against the reference's own known answers (test/sql/duckhts.test:736-758).  The reference's UDFs are
DuckDB scalar functions; harness/fake_duckdb hosts table functions only, so they cannot be run here: parity of the GPU columns is
this restatement applied to the CIGAR column, which is itself compared with the reference extension cell by cell."""

METRICS = {"S": "has_soft_clip", "H": "has_hard_clip", "l": "left_soft_clip", "r": "right_soft_clip", "q": "query_length",
           "a": "aligned_query_length", "R": "reference_length"}


def parse_cigar_metrics(cigar):
    """kmer_udf.c:197-269.  cigar: bytes.  Returns a dict of the seven metrics, or None where the function returns 0 (-> NULL)."""
    if cigar is None or len(cigar) == 0 or cigar == b"*":
        return None
    m = dict(has_soft_clip=0, has_hard_clip=0, left_soft_clip=0, right_soft_clip=0, query_length=0, aligned_query_length=0, reference_length=0)
    op_len, saw, first, last = 0, False, (None, 0), (None, 0)
    for c in cigar:
        ch = chr(c)
        if ch.isdigit() and c < 128:
            op_len = op_len * 10 + (c - 48)
            continue
        if op_len <= 0:
            return None
        if ch in "M=X":
            m["query_length"] += op_len; m["aligned_query_length"] += op_len; m["reference_length"] += op_len
        elif ch == "I":
            m["query_length"] += op_len
        elif ch == "S":
            m["query_length"] += op_len; m["has_soft_clip"] = 1
        elif ch == "H":
            m["has_hard_clip"] = 1
        elif ch in "DN":
            m["reference_length"] += op_len
        elif ch == "P":
            pass
        else:
            return None
        if not saw:
            first = (ch, op_len)
        last = (ch, op_len)
        saw, op_len = True, 0
    if not saw or op_len != 0:
        return None
    if first[0] == "S":
        m["left_soft_clip"] = first[1]
    if last[0] == "S":
        m["right_soft_clip"] = last[1]
    return m


def cigar_metric(cigar, metric):
    """cigar_metric_scalar (:695-742): metric = one of METRICS' keys.  Booleans come back as 0 / 1."""
    m = parse_cigar_metrics(cigar)
    return None if m is None else int(m[METRICS[metric]])


def cigar_has_op(cigar, op):
    """cigar_has_op_scalar (:744-790) over cigar_has_operator_text (:271-293).  op: one character (str)."""
    if len(op) != 1 or op.upper() not in "MIDNSHP=X":
        return None
    op = op.upper()
    if cigar is None or len(cigar) == 0 or cigar == b"*":
        return 0
    op_len = 0
    for c in cigar:
        ch = chr(c)
        if ch.isdigit() and c < 128:
            op_len = op_len * 10 + (c - 48)
            continue
        if op_len <= 0:
            return None
        if ch == op:
            return 1
        op_len = 0
    return None if op_len != 0 else 0
