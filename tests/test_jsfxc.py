"""CPU tests of the DSP-JSFX -> CUDA translator (za_gpu/jsfxc.py): the generated translation units compiled as
plain C++ (tools/emu_jsfx.py, -DZA_JSFX_HOST_EMU: same generated code, host stand-ins for the device intrinsics)
must reproduce the reference's compiled DSP (oracle/_ref) bit for bit, final variables included."""
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO / "tools"))
import za_oracle as zo


@pytest.mark.parametrize("key,auto", [("RTT", True), ("SOMA", False), ("ATTACK", True), ("ERBTilt", False),
                                      ("SpectralStabilizer", True), ("DDT", False), ("DPT", True), ("SaliencePush", False),
                                      ("ADS", True), ("BedRock", False), ("Roomalizer", True), ("TSEQ", False)])
def test_translation_is_bit_identical_on_host(key, auto):
    if not zo.have_ref(key):
        pytest.skip("oracle/_ref not built")
    import emu_jsfx
    err, bad = emu_jsfx.compare(key, nf=9000, automate=auto, verbose=False)
    assert err == 0.0, err
    assert not bad, bad[:4]


def test_parser_precedence_and_forms():
    from za_gpu import jsfxc
    P = lambda s: jsfxc.Parser(s).program()
    assert P("-x^2")[0] == ("bin", "^", ("un", "-", ("var", "x")), ("num", 2.0))          # prefix binds tighter than ^
    assert P("a ? b")[0] == ("tern", ("var", "a"), ("var", "b"), ("num", 0.0))            # missing ':' -> 0
    assert P("a = b = 1")[0][0] == "assign" and P("a = b = 1")[0][3][0] == "assign"       # right associative
    assert P("a\n  || b")[0] == ("bin", "||", ("var", "a"), ("var", "b"))                 # newline + infix continues
    assert len(P("a\n-b")) == 2                                                            # '-' may start a statement
    assert P("x & 3 == 3")[0][1] == "=="  or P("x & 3 == 3")[0][1] == "&"                 # same level, left to right
    assert P("loop(3, a += 1; b += 2)")[0][0] == "loop"
    with pytest.raises(NotImplementedError):
        jsfxc.Unit("t", "@init\nfunction f() instance(s) ( s = 1; );\n")
