"""Pins the CPU oracle against every known-answer test the reference holds for this path
(SURVEY.md 8(c)).  Inputs/expectations are the reference's own test vectors:
  COMMON_T/alignment/AlignmentTest.java:34-40
  COMMON_T/alignment/CigarStringUtilTest.java:34-120
  HSQ_T/process/FastqReadTrimmerTest.java:35-40
"""
import pytest

from oracle import oracle as O

T2_REF = "CATTCCAGGCCTAAGATCCCGTCCATCGCCACTGGGATGGTGGGGGCCCTCCTCTTGCTGCTGGTGGTGGCCCTGGGGATCGGCCTCTTCATGCGAAGGCGCCACATCGTTCGGAAGCGCACGCTGCGGAGGCTGCTGCAGGAGAGGGAG"
T2_QRY = "GTCCATCGCCAGTGGGATGGTGGGGGCCCTCCTCTTGCTGCTGGTGGTGGCCCTGGGGATCGGCCTCTTCATGCGAAGGCGCCACATCGTTCGGAAGCGCACGCTGCGGAGGCTGCTGCAGGAGAGGGAG"
T3_REF = "GACCCTTGTCTCTGTGTTCTTGTCCCCCCCAGCTTGTGGAGCCTCTTACACCCAGTGGAGAAGCTCCCAACCAAGCTCTCTTGAGGATCTTGAAGGAAACTGAATTCAAAAAGATCAAAGTGCTGGGCTCCGGTGCGTTCGGCACGGTGT"
T3_QRY = "GCCCCCCCAGCTTGTGGAGCCTCTTACACCCAGTGGAGAAGCTCCCAACCAAGCTCTCTTGAGGATCTTGAAGGAAACTGAATTCAAAAAGATCAAAGTGCTGGGCTCCGGTGCGTTCGGCACGGTGT"
T4_REF = "GACCCTTGTCTCTGTGTTCTTGTCCCCCCCAGCTTGTGGAGCCTCTTACACCCAGTGGAGAAGCTCCCAACCAAGCTCTCTTGAGGAGGGTCTTGAAGGAAACTGAATTCAAAAAGATCAAAGTGCTGGGCTCCGGTGCGTTCGGCACGGTGT"

# (reference, query, expected MD of the full alignment pair)   CigarStringUtilTest.java:34-72
MD_KATS = [
    ("AAAAAAAAAAATTTTT--GTTTTT", "AAAAAAAAAAGTTTTTACATTTTT", "10A5G5"),
    ("AAAAAAAAAAATTTTT--GCTTTTTTTTTTTTTT", "AAAAAAAAAAGTTTTTACAGTTTTTTTTTTTTTT", "10A5G0C14"),
    ("AAAAAAAAAAGTTTTTACATTTTT", "AAAAAAAAAAATTTTT--GTTTTT", "10G5^AC0A5"),
    (T2_REF, T2_QRY, "0^CATTCCAGGCCTAAGATCCC11C118"),
]


@pytest.mark.parametrize("ref,qry,md", MD_KATS)
def test_md_kats(ref, qry, md):
    out = O.align_batch([ref], [qry])
    assert out.mds()[0] == md


def test_md_three_trimmed_alignment_untrimmed_cigar():  # CigarStringUtilTest.java:75-84
    assert O.md_trimmed_with_full_cigar(T3_REF, T3_QRY) == "0^TCCCCCCCAGCTTGTGGAGCCT0C127"


def test_md_four():  # :87-96
    assert O.md_trimmed_with_full_cigar(T4_REF, T3_QRY) == "0^TCCCCCCCAGCTTGTGGAGCCT0C64^CAA63"


def test_md_reverse():  # :99-109
    assert O.md_trimmed_with_full_cigar(T2_REF[::-1], T2_QRY[::-1], reverse=True) == "0^GTCCATCGCCACTGGGATGG11C118"


def test_edit_distance():  # :112-120
    assert O.cigar_edit_distance("CATTCCG", "CACTCCG") == 1


def test_index_of_first_match_in_reference():  # AlignmentTest.java:34-40
    assert O.index_of_first_match_in_reference("ACTGAAATTTGGGCCCAGTCTAAANNNNTTTTT", "ATTTGGGC") == 6


def test_trim_kat():  # FastqReadTrimmerTest.java:35-40: 20 bp read, keep [5,14] -> 10 bases
    seq, qual = O.trim_batch(["ACGTACGTACGTACGTACGT"], ["ABCDEFGHIJKLMNOPQRST"], 5, 14, True)
    assert seq.strings() == ["CGTACGTACG"] and qual.strings() == ["FGHIJKLMNO"]
    seq, _ = O.trim_batch(["ACGTACGTACGTACGTACGT"], ["ABCDEFGHIJKLMNOPQRST"], 5, 14, False)
    assert seq.strings() == ["CGTACGTACGTACGT"]
    with pytest.raises(O.OracleError):
        O.trim_batch(["ACGT"], ["IIII"], 4, 10, True)


def test_find_counters_on_the_query_trimmed_alignment():
    """FastqFindAndTrimTool.java:930-938: counters over getAlignmentWithoutEndingAndBeginningQueryInserts
    (AlignmentPair.java:95-99,190-256), hand-derived from the alignment strings the oracle returns."""
    refs = ["TTTTACGTACGGTACGTAAAA", "ACGTACGTACGT", "ACGTACGT"]
    qrys = ["ACGTACGTTACGA", "ACGTTTACGTAGGT", "ACGTACGT"]
    a = O.align_batch(refs, qrys)
    for k in range(3):
        ra, qa = a.ref_alns()[k], a.qry_alns()[k]
        lo, hi = int(a.first_q[k]), int(a.last_q[k])
        ra, qa = ra[lo:hi + 1], qa[lo:hi + 1]
        ins_bases = ra.count("_")
        del_bases = qa.count("_")
        ins_gaps = len([g for g in ra.replace("A", " ").replace("C", " ").replace("G", " ").replace("T", " ").split() if g])
        del_gaps = len([g for g in qa.replace("A", " ").replace("C", " ").replace("G", " ").replace("T", " ").split() if g])
        mism = sum(1 for x, y in zip(ra, qa) if x != "_" and y != "_" and x != y)
        assert a.find_counts[k].tolist() == [mism, ins_gaps, ins_bases, del_gaps, del_bases], (k, ra, qa)
    assert a.find_counts[2].tolist() == [0, 0, 0, 0, 0]
    assert a.find_counts[:2].sum() > 0
