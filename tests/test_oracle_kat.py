"""Pin the CPU oracle against the reference's own known-answer tests (gen1/test.zig), machine-
transcribed into tests/golden/gen1_kat_cases.py by tools/transcribe_kat.py."""
import pytest

import gen1_kat_cases as kat
from kat_harness import KatFailure, SkipCase

RUNNABLE = [(n, ln, fn) for (n, ln, fn, why) in kat.CASES if fn is not None]
UNTRANSLATED = [(n, ln, why) for (n, ln, fn, why) in kat.CASES if fn is None]


@pytest.mark.parametrize("name,line,fn", RUNNABLE, ids=[f"{ln}:{n}" for n, ln, _ in RUNNABLE])
def test_reference_known_answer(name, line, fn):
    try:
        fn()
    except SkipCase:
        pytest.skip("reference skips this test in showdown mode")


def test_transcription_coverage():
    # the transcriber must keep translating the bulk of the reference suite
    assert len(RUNNABLE) >= 100, (len(RUNNABLE), UNTRANSLATED)
