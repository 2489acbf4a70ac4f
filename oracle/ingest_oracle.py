"""CPU restatement of the reference's ingest step — TEST INFRASTRUCTURE ONLY (tests/, never the product path).

Follows read_sequence (include/utils/io.hpp:6-18), the FASTA extraction operator it uses
(include/biovoltron/file_io/fasta.hpp:116-151), Codec::to_int (include/biovoltron/utility/istring.hpp:27-46) and the
`% 4` of suffix_sort_main (include/command/suffix_sort.hpp:33).

Pinned: tests/test_ingest.py compares it with the reference itself (oracle/_ref/libkiss_ref.so: ref_read_sequence =
the unmodified read_sequence on an istringstream) on every fixture and with tests/golden/ingest_golden.json, which
tests/golden/make_ingest_golden.py generated from the reference.
"""
import numpy as np

_LUT = np.zeros(256, np.uint8)          # Codec::ints (istring.hpp:28-36): 4 everywhere ... then % 4 -> 0
for _c, _v in ((b"aA", 0), (b"cC", 1), (b"gG", 2), (b"tT", 3)):
    for _b in _c:
        _LUT[_b] = _v


def _getlines(data):
    """what successive std::getline calls return: split at '\\n', delimiter dropped, no line after a final '\\n'"""
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


def read_sequence(data):
    """file bytes -> codes 0..3 (uint8).  FASTA mode iff the first byte is '>' (io.hpp:8)."""
    data = bytes(data)
    fasta = data[:1] == b">"
    keep = []
    prev_header = False
    for line in _getlines(data):
        # fasta.hpp:116-151: a record = its header line (:127), then lines are appended to the sequence
        # UNCONDITIONALLY (:129-133) until the byte after a line is '>' (:137-138).  So a line whose first byte is '>'
        # is a header only when the line before it was not one: of two header-looking lines in a row the second is
        # SEQUENCE ('>' and the letters encode like any byte), of three the third is a header again.
        # io.hpp:13-15: text mode keeps every line.
        header = fasta and line[:1] == b">" and not prev_header
        prev_header = header
        if not header:
            keep.append(line)            # '\r' of a CRLF file stays in the line (getline only strips '\n')
    seq = np.frombuffer(b"".join(keep), dtype=np.uint8)
    return _LUT[seq]
