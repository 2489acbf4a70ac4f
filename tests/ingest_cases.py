"""File-byte fixtures for the ingest step (shared by the CPU and GPU tests and the golden generator)."""
import numpy as np

import textgen as tg


def _fasta(seed, records, width, n_per, crlf=False, lower=False, with_n=False):
    rng = np.random.default_rng(seed)
    nl = b"\r\n" if crlf else b"\n"
    out = []
    for r in range(records):
        out.append(b">chr%d some description with > inside %d" % (r + 1, seed) + nl)
        L = int(rng.integers(max(1, n_per // 2), n_per + 1))
        s = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, L)].copy()
        if with_n and L > 40:
            p = int(rng.integers(0, L - 30))
            s[p:p + 25] = ord("N")
        if lower:
            s[::3] |= 0x20
        raw = s.tobytes()
        for i in range(0, L, width):
            out.append(raw[i:i + width] + nl)
    return b"".join(out)


CASES = {
    "empty": b"",
    "only_newline": b"\n",
    "header_only": b">chr1\n",
    "header_no_newline": b">chr1",
    "one_record": b">chr1 test\nACGTACGT\nTTGA\n",
    "no_trailing_newline": b">chr1\nACGT\nGG",
    "two_records": b">a\nACGT\n>b desc\nGGCC\nTT\n",
    "blank_lines": b">a\n\nAC\n\n\nGT\n>b\n\nTT\n\n",
    "gt_inside_line": b">a\nAC>GT\nA>\n",
    "lowercase_and_n": b">a\nacgtnNRYacgt\nNNNN\n",
    "crlf": b">a\r\nACGT\r\nGG\r\n>b\r\nTT\r\n",
    "text_mode": b"ACGT\nGGCC\n\nTTAA",
    "text_mode_gt_line": b"ACGT\n>not a header in text mode\nGG\n",
    "leading_newline_then_gt": b"\n>a\nACGT\n",          # first byte is not '>': text mode, the '>' line is sequence
    "spaces_and_tabs": b">a\nAC GT\n\tACGT \n",
    "long_header": b">" + b"x" * 9000 + b"\nACGTACGT\n>" + b"y" * 5000 + b"\nGGGG\n",
    "line_4096": b">a\n" + b"ACGT" * 1024 + b"\n" + b"TTGA" * 1024 + b"\n",
    # header-looking lines in a row: the line after a header is always sequence (fasta.hpp:129-133), so they alternate
    "two_headers_in_a_row": b">a\n>b\nACGT\n",
    "empty_record_in_the_middle": b">a\nAC\n>b\n>c\nGG\n",
    "three_headers_in_a_row": b">a\n>b\n>c\nACGT\n",
    "five_headers_in_a_row": b">a\n>bb\n>ccc\n>dddd\n>eeeee\nACGT\n>f\nTT\n",
    "headers_only_run": b">a\n>b\n>c\n>d",
    "blank_line_between_headers": b">a\n\n>b\n>c\nAC\n",
    "header_run_after_sequence": b">a\nACGT\n>b\n>c\n>d\n>e\nGG\n>f\n>g\n",
    "gt_only_lines": b">\n>\n>\n>\nA\n>\n>\n",
    # runs of header-looking lines across the 16-byte thread and 4096-byte chunk boundaries of the device kernels
    "header_run_across_threads": b">a\n" + b"".join(b">h%02d\n" % i for i in range(40)) + b"ACGT\n",
    "header_run_across_chunks": b">a\nAC\n" + b"".join(b">" + b"x" * (37 + i % 11) + b"\n" for i in range(400)) + b"GGTT\n",
    "long_header_run_odd": b">" + b"x" * 5000 + b"\n>" + b"y" * 5000 + b"\n>" + b"z" * 3000 + b"\nACGT\n>q\nTT\n",
    "long_header_run_even": b">" + b"x" * 5000 + b"\n>" + b"y" * 4100 + b"\nACGT\n>q\n>r\nTT\n",
    # (bytes >= 0x80 index the reference's 128-entry table out of range — undefined there, 0 here: no fixture)
}


_BIG = None


def big_cases():
    global _BIG
    if _BIG is None:
        _BIG = _big_cases()
    return _BIG


def _big_cases():
    text = np.frombuffer(b"ACGT", np.uint8)[tg.uniform(7, 300_000)].tobytes()
    return {
        "fasta_60": _fasta(1, 7, 60, 40_000),
        "fasta_80_crlf": _fasta(2, 5, 80, 30_000, crlf=True),
        "fasta_lower_n": _fasta(3, 9, 70, 25_000, lower=True, with_n=True),
        "fasta_wide_lines": _fasta(4, 3, 10_000, 120_000),
        "fasta_width1": _fasta(5, 4, 1, 3_000),
        "fasta_many_records": _fasta(6, 400, 60, 300),
        "text_big": b"\n".join(text[i:i + 100] for i in range(0, 300_000, 100)),
    }
