"""CSV inputs shared by the golden generator (tests/golden/make_csv_golden.py), the oracle tests and
the GPU parity tests. Everything is built deterministically from code, so that the large cases
(fields at the size limit, files spanning many 16 KiB scan blocks) need no fixture files.
"""
from __future__ import annotations

import random
from typing import Dict, List, Tuple

Query = Tuple[int, int, str]  # start_row, shard_size, mode

BASIC_QUERIES: List[Query] = [(0, 100, "rows"), (0, 1, "rows"), (1, 2, "rows"), (2, 1000, "count"),
                              (0, 3, "count"), (7, 5, "rows")]


def _fuzz(seed: int, length: int, alphabet: str) -> bytes:
    rng = random.Random(seed)
    return "".join(rng.choice(alphabet) for _ in range(length)).encode("utf-8")


def _table(rows: int, seed: int, newline: str = "\n", quote_rate: float = 0.3) -> bytes:
    """An `id,text,source` table like BASELINE config 5, with quoting, embedded commas, quotes and
    line ends in some text cells."""
    rng = random.Random(seed)
    words = ["alpha", "beta", "gamma", "delta", "naïve", "Zürich", "東京", "data", "set", "row", "😀", "x"]
    out = ["id,text,source"]
    for i in range(rows):
        k = rng.randint(1, 24)
        text = " ".join(rng.choice(words) for _ in range(k))
        r = rng.random()
        if r < quote_rate / 3:
            text = '"' + text.replace(" ", ", ", 2) + '"'
        elif r < 2 * quote_rate / 3:
            text = '"' + text.replace(" ", ' ""q"" ', 1) + '"'
        elif r < quote_rate:
            text = '"' + text.replace(" ", rng.choice(["\n", "\r\n", "\r"]), 1) + '"'
        out.append(f"{i},{text},src{rng.randint(0, 9)}")
    return (newline.join(out) + newline).encode("utf-8")


def small_cases() -> Dict[str, bytes]:
    c: Dict[str, bytes] = {}
    c["basic"] = b"id,text,source\n1,hello world,a\n2,second row,b\n3,third,c\n"
    c["crlf"] = b"id,text\r\n1,one\r\n2,two\r\n"
    c["cr_only"] = b"a,b\r1,2\r\r3,4,5\r\n6\n"
    c["mixed_newlines"] = b"a,b\n1,2\r\n3,4\r5,6\n\r7,8\r\n\n9,10"
    c["no_trailing_newline"] = b"a,b\n1,2\n3,4"
    c["trailing_blank_lines"] = b"a,b\n1,2\n\n\n\r\n\r"
    c["leading_blank_line"] = b"\n\na,b\n1,2\n"
    c["leading_crlf"] = b"\r\na,b\r\n1,2\r\n"
    c["leading_cr"] = b"\ra,b\r1,2\r"
    c["spaces_only_lines"] = b"a,b\n \n1,2\n  ,\n"
    c["quoted_commas"] = b'id,text\n1,"a, b, c"\n2,"x,y"\n'
    c["quoted_newlines"] = b'id,text,z\n1,"line one\nline two",q\n2,"crlf\r\ninside",r\n3,"cr\rinside",s\n'
    c["doubled_quotes"] = b'a,b\n"say ""hi""",2\n"""",""""""\n'
    c["quote_mid_field"] = b'a,b\nab"cd,e"f\n5" tall,"x"\n'
    c["text_after_close"] = b'a,b\n"ab"cd,"x"y"z"\n"p" ,q\n'
    c["quote_then_quote_text"] = b'a\n"ab""cd"ef"gh\n'
    c["unterminated_quote"] = b'a,b\n1,"never closed\nmore,stuff\n'
    c["unterminated_quote_empty"] = b'a,b\n1,"'
    c["lone_quote"] = b'"'
    c["two_quotes"] = b'""'
    c["quote_newline_eof"] = b'a\n"\n'
    c["empty_quoted_fields"] = b'a,b,c\n"","",""\n,,\n"",,x\n'
    c["extra_fields"] = b"a,b\n1,2,3,4\n5,6\n"
    c["missing_fields"] = b"a,b,c\n1\n2,3\n4,5,6\n"
    c["duplicate_names"] = b"a,a,b\n1,2,3\n"
    c["quoted_header"] = b'"id","the, text","so""urce"\n1,2,3\n'
    c["multiline_header"] = b'"id\npart two",text\n1,2\n'
    c["bom"] = b"\xef\xbb\xbfid,text\n1,x\n"
    c["non_ascii"] = "id,text\n1,naïve café\n2,東京タワー\n3,😀 emoji 😀\n4,\"quoted ü, with comma\"\n".encode()
    c["nul_byte"] = b"a,b\n1,\x002\n\x00,3\n"
    c["empty_file"] = b""
    c["header_only"] = b"a,b,c"
    c["header_only_newline"] = b"a,b,c\n"
    c["only_newlines"] = b"\n\n\r\n\r\n\n"
    c["single_newline"] = b"\n"
    c["single_cr"] = b"\r"
    c["lf_cr_pairs"] = b"a\n\rb\n\rc\n\r"
    c["comma_only"] = b",\n,\n,,\n"
    c["trailing_comma_eof"] = b"a,b\n1,"
    c["tabs_and_spaces"] = b"a,b\n\t1 , 2\t\n"
    c["quote_after_space"] = b'a,b\n "x",y\n'
    c["short_rows_dense"] = b"h\n" + b"a\n" * 300
    c["single_chars"] = b"h\n" + b"a\nb\r\nc\r" * 120
    # rows and quoted fields placed across the 64-byte chunk and 16 KiB block boundaries of the scan
    c["chunk_boundary_rows"] = b"h1,h2\n" + b"".join(
        (b"x" * k + b',"' + b"y" * (63 - k % 7) + b'\n"",z"\r\n') for k in range(1, 140))
    c["block_spanning_quote"] = b'id,text\n1,"' + b"long, \n" * 5000 + b'"\n2,short\n3,"' + b"q" * 16384 + b'"\n4,end\n'
    c["crlf_on_boundaries"] = b"a,b\r\n" + b"".join(b"r" * (62 - (i % 3)) + b",1\r\n" for i in range(600))
    c["quotes_on_boundaries"] = b"a\n" + b"".join(b'"' + b"w" * (60 + i % 5) + b'""' + b'"\n' for i in range(500))
    c["table_lf"] = _table(400, 1)
    c["table_crlf"] = _table(400, 2, "\r\n")
    c["table_cr"] = _table(200, 3, "\r")
    for seed in range(24):
        c[f"fuzz_struct_{seed}"] = _fuzz(1000 + seed, 700 + 37 * seed, 'ab,"\n\r ')
    for seed in range(8):
        c[f"fuzz_quotes_{seed}"] = _fuzz(2000 + seed, 3000, 'a,""\n')
    for seed in range(8):
        c[f"fuzz_unicode_{seed}"] = _fuzz(3000 + seed, 1500, 'aé東😀,"\n\r')
    return c


def error_cases() -> Dict[str, Tuple[bytes, List[Query]]]:
    """Inputs the reference fails on — but only when it reads far enough, hence the queries."""
    e: Dict[str, Tuple[bytes, List[Query]]] = {}
    row = b"1234567,abcdefghijklmnopqrstuvwxyz\n"  # 35 bytes
    head = b"id,text\n"

    def with_bad(at: int, bad: bytes, total_rows: int = 1200) -> bytes:
        body = bytearray(head + row * total_rows)
        body[at:at + len(bad)] = bad
        return bytes(body)

    qs: List[Query] = [(0, 1, "rows"), (0, 100, "count"), (150, 50, "rows"), (225, 8, "rows"), (230, 5, "count"),
                       (400, 100, "rows"), (460, 10, "count"), (0, 5000, "count"), (1190, 20, "rows")]
    e["bad_start_byte_early"] = (with_bad(20, b"\xff"), qs)
    e["bad_start_byte_chunk1_end"] = (with_bad(8191, b"\x80"), qs)
    e["bad_start_byte_chunk2_start"] = (with_bad(8192, b"\xfe"), qs)
    e["bad_start_byte_chunk2_mid"] = (with_bad(12000, b"\xc0"), qs)
    e["bad_continuation_mid"] = (with_bad(16000, b"\xe2\x28\xa1"), qs)
    e["sequence_across_reads_ok"] = (with_bad(8191, "é".encode()), qs)
    e["sequence_across_reads_bad"] = (with_bad(8191, b"\xc3\x28"), qs)
    e["three_byte_across_reads_bad"] = (with_bad(16382, b"\xe6\x9d\x28"), qs)
    e["overlong"] = (with_bad(9000, b"\xe0\x80\xaf"), qs)
    e["surrogate"] = (with_bad(9000, b"\xed\xa0\x80"), qs)
    e["too_large"] = (with_bad(9000, b"\xf4\x90\x80\x80"), qs)
    e["f5"] = (with_bad(30000, b"\xf5"), qs)
    e["truncated_at_eof"] = (head + row * 300 + b"9,abc\xe6\x9d", qs)
    e["truncated_at_eof_1"] = (head + row * 300 + b"9,abc\xf0", qs)
    e["bad_in_header"] = (b"id,te\xffxt\n" + row * 10, [(0, 1, "rows"), (0, 100, "count")])
    e["bad_last_byte"] = (head + row * 300 + b"\x80", qs)

    limit = 131072
    big_ok = b"k," + b"x" * limit + b"\n"
    big_bad = b"k," + b"x" * (limit + 1) + b"\n"
    big_bad_quoted = b'k,"' + b'y""' * 50000 + b"z" * (limit - 100000 + 1) + b'"\n'
    big_ok_multibyte = b"k," + "é".encode() * limit + b"\n"
    big_bad_multibyte = b"k," + "東".encode() * (limit + 1) + b"\n"
    lq: List[Query] = [(0, 1, "rows"), (0, 2, "count"), (0, 3, "count"), (0, 4, "count"), (2, 1, "count"),
                       (3, 1, "count"), (3, 5, "rows"), (10, 5, "count")]
    few = row * 3
    e["field_at_limit"] = (head + few + big_ok + few, lq)
    e["field_over_limit"] = (head + few + big_bad + few, lq)
    e["field_over_limit_quoted"] = (head + few + big_bad_quoted + few, lq)
    e["field_at_limit_multibyte"] = (head + few + big_ok_multibyte + few, lq)
    e["field_over_limit_multibyte"] = (head + few + big_bad_multibyte + few, lq)
    e["header_over_limit"] = (b"id," + b"h" * (limit + 1) + b"\n" + few, [(0, 1, "rows"), (5, 1, "count")])
    e["header_at_limit"] = (b"id," + b"h" * limit + b"\n" + few, [(0, 1, "count"), (5, 1, "count")])
    # both problems in one file, in either order
    e["bad_byte_then_long_field"] = (head + few + b"2,\xff\n" + big_bad + few, lq)
    e["long_field_then_bad_byte"] = (head + few + big_bad + few + b"2,\xff\n" + row * 400, lq)
    e["long_field_with_bad_byte_inside_line"] = (head + few + b"k," + b"x" * (limit + 50) + b"\xff" + b"x" * 9000 + b"\n" + few, lq)
    return e


def all_cases() -> Dict[str, Tuple[bytes, List[Query]]]:
    out: Dict[str, Tuple[bytes, List[Query]]] = {k: (v, BASIC_QUERIES) for k, v in small_cases().items()}
    out.update(error_cases())
    return out
