"""CPU tier, SURVEY.md §8 a6'(ii): the reference's table-space limits.  The vectors of tests/tablespace.py are
driven through the compiled reference, whose inflateCodesUsed() (src/inflate.c:1532) proves the entry counts
(852 / 850 literal/length entries, 592 / 594 distance entries) and whose verdicts pin the Deflate64-only limits
851 / 593 (src/inftrees.c:211-212,226-228,295-297).  The same streams are part of cases.probe_vectors(), so the
restatement, the host emulation of the device decoders and every GPU decoder configuration run them too."""
import pytest

import cases
import orclib
import tablespace

REF = orclib.ref()
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built (no /root/reference here)")


def test_count_vectors_are_complete_and_need_the_stated_space():
    for (nsym, root, used), count in tablespace.VECTORS.items():
        assert tablespace.kraft_complete(count) and sum(count) <= nsym
        assert tablespace.table_space(count, root) == used


@needs_ref
def test_reference_allocates_exactly_the_stated_entries():
    two = 2                                        # the companion code of every vector: two one-bit codes, 2 entries
    for name, data, wb, cap in tablespace.probe_cases():
        used = int(name.split("_")[3])
        entries, r = REF.codes_used(data, wb)
        if r.ret == 1:
            assert entries == used + two, (name, entries)
        else:
            assert entries in (0, two), (name, entries)     # rejected before (0) or after (2) the first table was kept


@needs_ref
def test_limits_differ_between_the_two_modes_exactly_where_the_reference_says():
    got = {(n, wb): REF.oneshot(d, cap, wb)[0] for n, d, wb, cap in tablespace.probe_cases()}
    def ok(key, wb):
        return got[("tabspace_%s_wb%d" % (key, wb), wb)].ret == 1
    assert ok("286sym_root9_852", -15) and not ok("286sym_root9_852", -16)
    assert ok("286sym_root9_850", -15) and ok("286sym_root9_850", -16)
    assert ok("32sym_root6_592", -16) and not ok("32sym_root6_594", -16)
    assert ok("30sym_root6_592", -15) and ok("30sym_root6_592", -16)
    assert got[("tabspace_286sym_root9_852_wb-16", -16)].msg_id == 12      # invalid literal/lengths set
    assert got[("tabspace_32sym_root6_594_wb-16", -16)].msg_id == 13       # invalid distances set


def test_vectors_are_in_the_shared_probe_list():
    names = {c[0] for c in cases.probe_vectors()}
    assert {c[0] for c in tablespace.probe_cases()} <= names
