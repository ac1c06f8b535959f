"""`wtns check` oracle on the SYNTH-1 circuit: a generated witness satisfies every constraint, a tampered one fails at the
first constraint that reads the changed wire.  CPU only."""
import pytest

from oracle import binfile, bn254 as bn, factory, wtns_check as WC, zkey_new as ZN


def test_oracle_accepts_the_generated_witness_and_finds_the_first_broken_constraint():
    circ = factory.SynthCircuit(n_vars=85, n_public=2, n_constraints=100, seed=0xB2000001)
    r1cs = ZN.write_r1cs(circ)
    w = circ.witness()
    assert WC.wtns_check(r1cs, binfile.write_wtns(w)) is None
    for wire in (1, 7, 40, 84):
        bad = list(w)
        bad[wire] = (bad[wire] + 2) % bn.R          # + 1 would turn a bit 0 into a bit 1: still a witness
        first = next(c for c in range(circ.n_constraints) if any(s == wire for lc in circ.constraint(c) for s, _ in lc))
        got = WC.wtns_check(r1cs, binfile.write_wtns(bad))
        assert got is not None and got >= first
        assert any(s == wire for lc in circ.constraint(got) for s, _ in lc)
    with pytest.raises(ValueError, match="Invalid witness length"):
        WC.wtns_check(r1cs, binfile.write_wtns(w[:-1]))
