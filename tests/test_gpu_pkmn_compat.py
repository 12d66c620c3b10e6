"""Config C1 through the libpkmn-compatible scalar symbols (include/oakb200_pkmn.h): single sample-team battles, fixed
seeds, played exactly as Oak's rollout loop calls libpkmn (cpp/include/search/mcts.h:448-496: choices x2, pick,
options_set(NULL...), update) and compared byte for byte with the oracle after every update: battle, durations,
actions, result, legal choices. Also the calc-override call pattern of the tree step (mcts.h:569-590)."""
import ctypes

import numpy as np
import pytest

import oak_b200
from oak_ref import Oracle, pick_word, sample_seed

pytestmark = pytest.mark.gpu


class Pkmn:
    def __init__(self):
        self.lib = lib = oak_b200.load_library()
        vp = ctypes.c_void_p
        lib.pkmn_gen1_battle_update.restype = ctypes.c_uint8
        lib.pkmn_gen1_battle_update.argtypes = [vp, ctypes.c_uint8, ctypes.c_uint8, vp]
        lib.pkmn_gen1_battle_choices.restype = ctypes.c_uint8
        lib.pkmn_gen1_battle_choices.argtypes = [vp, ctypes.c_uint8, ctypes.c_uint8, vp, ctypes.c_size_t]
        lib.pkmn_gen1_battle_options_set.restype = None
        lib.pkmn_gen1_battle_options_set.argtypes = [vp, vp, vp, vp]
        lib.pkmn_gen1_battle_options_chance_durations.restype = vp
        lib.pkmn_gen1_battle_options_chance_durations.argtypes = [vp]
        lib.pkmn_gen1_battle_options_chance_actions.restype = vp
        lib.pkmn_gen1_battle_options_chance_actions.argtypes = [vp]
        self.opts = np.zeros(128, np.uint8)

    def choices(self, battle, player, request):
        out = np.zeros(9, np.uint8)
        n = self.lib.pkmn_gen1_battle_choices(battle.ctypes.data, player, request, out.ctypes.data, 9)
        return [int(x) for x in out[:n]]

    def options_set(self, chance=None, calc=None):
        self.lib.pkmn_gen1_battle_options_set(self.opts.ctypes.data, None, None if chance is None else chance.ctypes.data,
                                              None if calc is None else calc.ctypes.data)

    def update(self, battle, c1, c2, with_options=True):
        return self.lib.pkmn_gen1_battle_update(battle.ctypes.data, c1, c2, self.opts.ctypes.data if with_options else None)

    def durations(self):
        return np.ctypeslib.as_array(ctypes.cast(self.lib.pkmn_gen1_battle_options_chance_durations(self.opts.ctypes.data),
                                                 ctypes.POINTER(ctypes.c_uint8)), (8,)).copy()

    def actions(self):
        return np.ctypeslib.as_array(ctypes.cast(self.lib.pkmn_gen1_battle_options_chance_actions(self.opts.ctypes.data),
                                                 ctypes.POINTER(ctypes.c_uint8)), (16,)).copy()


def test_single_battle_traces_are_bit_exact():
    orc, pk = Oracle(), Pkmn()
    total = 0
    for i in (0, 1, 37, 255):
        ob, od, orr = orc.sample_battle(i)
        gb, r = ob.copy(), orr
        chance = np.zeros(40, np.uint8)  # {probability 16, actions 16, durations 8}: install zero durations once
        pk.options_set(chance=chance)
        seed = sample_seed(i)
        for step in range(2000):
            if r & 15:
                break
            c = [pk.choices(gb, p, (r >> (4 + 2 * p)) & 3) for p in (0, 1)]
            assert c == [orc.choices(ob, p, (r >> (4 + 2 * p)) & 3) for p in (0, 1)]
            w = pick_word(seed, step)
            c1, c2 = c[0][w % len(c[0])], c[1][(w >> 32) % len(c[1])]
            pk.options_set()  # reset: keeps durations, clears actions (mcts.h:462)
            r = pk.update(gb, c1, c2)
            o_r, o_act = orc.update(ob, c1, c2, od)
            assert r == o_r and np.array_equal(gb, ob) and np.array_equal(pk.durations(), od) and np.array_equal(pk.actions(), o_act)
            total += 1
        assert r & 15
    assert total > 200


def test_null_options_and_calc_overrides():
    orc, pk = Oracle(), Pkmn()
    ob, od, r = orc.sample_battle(5)
    gb = ob.copy()
    for step in range(40):
        if r & 15:
            break
        c = [orc.choices(ob, p, (r >> (4 + 2 * p)) & 3) for p in (0, 1)]
        c1, c2 = c[0][step % len(c[0])], c[1][(step * 7) % len(c[1])]
        if step % 2:
            r2 = pk.update(gb, c1, c2, with_options=False)  # NULL options: state still advances identically
            r, _ = orc.update(ob, c1, c2, od)
        else:
            calc = np.zeros(24, np.uint8)
            calc[0], calc[8] = 217 + step % 39, 255 - step % 39  # the damage-roll override bytes Oak sets (mcts.h:569-590)
            pk.options_set(calc=calc)
            r2 = pk.update(gb, c1, c2)
            r, _ = orc.update(ob, c1, c2, od, overrides=calc[:16])
        assert r2 == r and np.array_equal(gb, ob)


def test_choices_truncation_and_pass_request():
    """pkmn.h:270-288: at most `len` entries are written, the return value is the number of choices there are; a PASS request
    has the single choice 0."""
    orc, pk = Oracle(), Pkmn()
    b, d, r = orc.sample_battle(2)
    full = orc.choices(b, 0, 1)
    assert len(full) >= 5
    out = np.full(9, 0xEE, np.uint8)
    n = pk.lib.pkmn_gen1_battle_choices(b.ctypes.data, 0, 1, out.ctypes.data, 3)
    assert n == len(full) and out[:3].tolist() == full[:3] and np.all(out[3:] == 0xEE)
    for player in (0, 1):
        assert pk.choices(b, player, 0) == [0] == orc.choices(b, player, 0)
    assert pk.lib.pkmn_gen1_battle_choices(b.ctypes.data, 2, 1, out.ctypes.data, 9) == 0  # not a player
    assert pk.lib.pkmn_gen1_battle_choices(b.ctypes.data, 0, 1, out.ctypes.data, 0) == 0  # no room at all


def test_scalar_calls_from_many_threads_do_not_interfere():
    """One context per calling thread: eight threads each play their own battle through the scalar symbols concurrently."""
    import threading
    orc = Oracle()
    errors = []

    def play(i):
        try:
            pk = Pkmn()
            local = Oracle()
            ob, od, r = local.sample_battle(i)
            gb = ob.copy()
            pk.options_set(chance=np.zeros(40, np.uint8))
            seed = sample_seed(i)
            for step in range(60):
                if r & 15:
                    break
                c = [local.choices(ob, p, (r >> (4 + 2 * p)) & 3) for p in (0, 1)]
                w = pick_word(seed, step)
                c1, c2 = c[0][w % len(c[0])], c[1][(w >> 32) % len(c[1])]
                pk.options_set()
                r2 = pk.update(gb, c1, c2)
                r, _ = local.update(ob, c1, c2, od)
                assert r2 == r and np.array_equal(gb, ob) and np.array_equal(pk.durations(), od), (i, step)
        except Exception as e:  # noqa: BLE001
            errors.append((i, repr(e)))

    threads = [threading.Thread(target=play, args=(i,)) for i in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
