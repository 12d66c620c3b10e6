#!/usr/bin/env python3
"""Generate the Gen 1 constant tables from the reference's generated Zig data.

Reads (never copies verbatim) /root/reference/cpp/extern/engine/src/lib/gen1/data/
{moves,species,types}.zig and cross-checks against engine/src/data/data.json, then emits

  oracle/gen1_tables.h            plain-C arrays for the CPU oracle
  oak_b200/csrc/gen1_tables.cuh   the same numbers for the CUDA engine (constant memory)
  tests/golden/gen1_names.json    enum name -> ordinal maps (used by the KAT transcriber)

Reference: moves.zig:12-181 (Move enum), :196-1517 (DATA), :1520-1596 (Effect enum),
:1663-1829 (PP); species.zig:11-163 (enum), :165-318 (CHANCES), :327-1084 (DATA);
types.zig:15-48 (Type enum, CHART).
Run only in the build container (the reference tree does not exist on the GPU box); the
generated files are committed.
"""
import json
import os
import re
import sys

REF = "/root/reference/cpp/extern/engine/src"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def enum_names(text, header):
    """Names of a `pub const X = enum(..) {` block up to the first blank line / const."""
    start = text.index(header)
    body = text[start + len(header):]
    names = []
    for line in body.splitlines():
        s = line.strip()
        if not s:
            if names:
                break
            continue
        if s.startswith("//"):
            continue
        m = re.match(r"^([A-Za-z_][A-Za-z0-9_]*),$", s)
        if m:
            names.append(m.group(1))
        else:
            if names:
                break
    return names


def main():
    moves_zig = open(f"{REF}/lib/gen1/data/moves.zig").read()
    species_zig = open(f"{REF}/lib/gen1/data/species.zig").read()
    types_zig = open(f"{REF}/lib/gen1/data/types.zig").read()
    data_json = json.load(open(f"{REF}/data/data.json"))[0]

    types = enum_names(types_zig, "pub const Type = enum(u4) {")
    assert types == data_json["types"], "types.zig vs data.json mismatch"
    tid = {n: i for i, n in enumerate(types)}

    moves = enum_names(moves_zig, "pub const Move = enum(u8) {")
    assert moves[0] == "None" and len(moves) == 166, len(moves)
    effects = enum_names(moves_zig, "pub const Effect = enum(u8) {")
    # the Effect enum has section comments between groups; re-scan permissively
    eff_block = moves_zig[moves_zig.index("pub const Effect = enum(u8) {"):]
    eff_block = eff_block[:eff_block.index("comptime")]
    effects = re.findall(r"^\s+([A-Za-z0-9]+),$", eff_block, re.M)
    assert effects[0] == "None" and effects[-1] == "OHKO" and len(effects) == 68, len(effects)
    eid = {n: i for i, n in enumerate(effects)}

    # Move.DATA
    data_block = moves_zig[moves_zig.index("const DATA = [_]Data{"):moves_zig.index("/// Representation of a move's effect.")]
    entries = re.findall(
        r"//\s*(\w+)\s*\n\s*\.\{\s*\.effect = \.(\w+),\s*\.bp = (\d+),\s*\.type = \.(\w+),\s*"
        r"\.accuracy = percent\((\d+)\),\s*\.target = \.(\w+),\s*\}", data_block)
    assert len(entries) == 165, len(entries)
    move_data = []
    for i, (name, eff, bp, typ, acc, target) in enumerate(entries):
        assert name == moves[i + 1], (name, moves[i + 1])
        move_data.append(dict(name=name, effect=eid[eff], bp=int(bp), type=tid[typ],
                              accuracy=(int(acc) * 0xFF) // 100, self=int(target == "Self")))

    pp_block = moves_zig[moves_zig.index("const PP = [_]u8{"):]
    pp_block = pp_block[:pp_block.index("};")]
    pps = re.findall(r"^\s+(\d+), // (\w+),?$", pp_block, re.M)
    assert len(pps) == 165, len(pps)
    json_pp = {re.sub(r"[^a-z0-9]", "", k.lower()): v for k, v in data_json["moves"].items()}
    for i, (pp, name) in enumerate(pps):
        assert name == moves[i + 1]
        assert json_pp[name.lower()] == int(pp), name
    pp = [int(p) for p, _ in pps]

    species = enum_names(species_zig, "pub const Species = enum(u8) {")
    assert species[0] == "None" and len(species) == 152
    ch_block = species_zig[species_zig.index("const CHANCES = [_]u8{"):]
    ch_block = ch_block[:ch_block.index("};")]
    chances = re.findall(r"^\s+(\d+), // (\w+)$", ch_block, re.M)
    assert len(chances) == 151
    sp_block = species_zig[species_zig.index("const DATA = [_]Data{"):]
    sp_entries = re.findall(
        r"//\s*(\w+)\s*\n\s*\.\{\s*\.stats = \.\{ \.hp = (\d+), \.atk = (\d+), \.def = (\d+), "
        r"\.spe = (\d+), \.spc = (\d+) \},\s*\.types = \.\{ \.type1 = \.(\w+), \.type2 = \.(\w+) \},",
        sp_block)
    assert len(sp_entries) == 151, len(sp_entries)
    species_data = []
    json_sp = {re.sub(r"[^a-z0-9]", "", k.lower()): v for k, v in data_json["species"].items()}
    for i, (name, hp, atk, de, spe, spc, t1, t2) in enumerate(sp_entries):
        assert name == species[i + 1] == chances[i][1]
        j = json_sp[name.lower()]
        assert [j["stats"][k] for k in ("hp", "atk", "def", "spe", "spc")] == [int(hp), int(atk), int(de), int(spe), int(spc)]
        jt = j["types"] if len(j["types"]) == 2 else j["types"] * 2
        assert jt == [t1, t2], name
        # Species.chance: base speed / 2 (species.zig CHANCES)
        assert int(chances[i][0]) == int(spe) // 2, name
        species_data.append(dict(name=name, stats=[int(hp), int(atk), int(de), int(spe), int(spc)],
                                 types=tid[t1] | (tid[t2] << 4), chance=int(chances[i][0])))

    chart_block = types_zig[types_zig.index("const CHART"):types_zig.index("const PRECEDENCE")]
    rows = re.findall(r"\[_\]Effectiveness\{([^}]*)\}", chart_block)
    assert len(rows) == 15
    val = {"S": 20, "N": 10, "R": 5, "I": 0}
    chart = [[val[c.strip()] for c in r.split(",")] for r in rows]
    assert all(len(r) == 15 for r in chart)

    def arr(name, ctype, vals, per_line=16):
        out = [f"{ctype} {name}[{len(vals)}] = {{"]
        for i in range(0, len(vals), per_line):
            out.append("    " + ", ".join(str(v) for v in vals[i:i + per_line]) + ",")
        out.append("};")
        return "\n".join(out)

    # Move.Data packed as u32: effect | bp<<8 | accuracy<<16 | type<<24 | self<<28
    # (our own packing; index 0 is an all-zero pad so tables are indexed by move id directly)
    packed = [0] + [m["effect"] | (m["bp"] << 8) | (m["accuracy"] << 16) | (m["type"] << 24) | (m["self"] << 28)
                    for m in move_data]
    pp0 = [0] + pp
    chance0 = [0] + [s["chance"] for s in species_data]
    sp_stats = [0] * 5 + [v for s in species_data for v in s["stats"]]
    sp_types = [0] + [s["types"] for s in species_data]
    chart_flat = [v for r in chart for v in r]

    hdr = ("// GENERATED by tools/gen_tables.py from the reference's generated data tables\n"
           "// (engine/src/lib/gen1/data/{moves,species,types}.zig, cross-checked with engine/src/data/data.json).\n"
           "// Do not edit. Move data word = effect | bp<<8 | accuracy<<16 | type<<24 | target_is_self<<28.\n")

    def body(q):
        s = []
        s.append(arr("GEN1_MOVE_DATA", q + " unsigned int", packed, 8))
        s.append(arr("GEN1_MOVE_PP", q + " unsigned char", pp0))
        s.append(arr("GEN1_SPECIES_CRIT", q + " unsigned char", chance0))
        s.append(arr("GEN1_SPECIES_STATS", q + " unsigned char", sp_stats, 15))
        s.append(arr("GEN1_SPECIES_TYPES", q + " unsigned char", sp_types))
        s.append(arr("GEN1_TYPE_CHART", q + " unsigned char", chart_flat, 15))
        return "\n\n".join(s) + "\n"

    enums = ["enum Gen1Effect {"] + [f"    EFF_{n} = {i}," for i, n in enumerate(effects)] + ["};"]
    mv_enum = ["enum Gen1Move {"] + [f"    MOVE_{n} = {i}," for i, n in enumerate(moves)] + ["    MOVE_SKIP_TURN = 255,", "};"]
    ty_enum = ["enum Gen1Type {"] + [f"    TYPE_{n} = {i}," for i, n in enumerate(types)] + ["};"]
    enum_txt = "\n".join(enums) + "\n\n" + "\n".join(mv_enum) + "\n\n" + "\n".join(ty_enum) + "\n"

    # protocol-log ordinals (common/protocol.zig:16-311): ArgType and the reason enums of the Generation I messages
    proto_src = open(f"{REF}/lib/common/protocol.zig").read()
    protocol = {}
    for name in ("ArgType", "Move", "Cant", "Heal", "Damage", "Status", "CureStatus", "Boost", "Fail", "Activate", "Start", "End",
                 "Immune"):
        m = re.search(r"pub const " + name + r" = enum\(u8\) \{(.*?)\n\};", proto_src, re.S)
        body_ = re.sub(r"//[^\n]*", "", m.group(1))
        protocol[name] = [w.strip() for w in body_.split(",") if w.strip()]
    assert protocol["ArgType"][30] == "Transform" and protocol["End"][22] == "ReflectSilent"
    GEN1_ARGS = protocol["ArgType"][:31]
    KEEP = {"Cant": 7, "Heal": 3, "Damage": 6, "Status": 3, "CureStatus": 2, "Boost": 8, "Fail": 9, "Activate": 7, "Start": 12,
            "End": 24, "Immune": 2}
    PREFIX = {"Cant": "CANT", "Heal": "HEAL", "Damage": "DMG", "Status": "STS", "CureStatus": "CURE", "Boost": "BST", "Fail": "FAIL",
              "Activate": "ACT", "Start": "START", "End": "END", "Immune": "IMM"}
    lg = ["enum LogArg {"] + [f"    LG_{n} = {i}," for i, n in enumerate(GEN1_ARGS)] + ["};"]
    for k, n_keep in KEEP.items():
        lg += [f"enum Log{k} {{"] + [f"    {PREFIX[k]}_{n} = {i}," for i, n in enumerate(protocol[k][:n_keep])] + ["};"]
    log_txt = "\n// protocol-log ordinals (reference engine/src/lib/common/protocol.zig:16-311)\n" + "\n".join(lg) + "\n"

    os.makedirs(f"{ROOT}/oracle", exist_ok=True)
    os.makedirs(f"{ROOT}/oak_b200/csrc", exist_ok=True)
    os.makedirs(f"{ROOT}/tests/golden", exist_ok=True)
    with open(f"{ROOT}/oracle/gen1_tables.h", "w") as f:
        f.write(hdr + "#pragma once\n\n" + enum_txt + "\n" + body("static const"))
    with open(f"{ROOT}/oak_b200/csrc/gen1_tables.cuh", "w") as f:
        f.write(hdr + "#pragma once\n\n" + enum_txt.replace("Gen1Effect", "Effect").replace("Gen1Move", "Move").replace("Gen1Type", "Type") + log_txt
                + "\n#ifdef __CUDACC__\n#define OAK_TABLE static __device__ const\n#else\n#define OAK_TABLE static const\n#endif\n\n"
                + body("OAK_TABLE"))
    with open(f"{ROOT}/tests/golden/gen1_names.json", "w") as f:
        json.dump(dict(moves=moves, species=species, types=types, effects=effects,
                       move_data=move_data, pp=pp, species_data=species_data, chart=chart, protocol=protocol), f)
    print(f"moves={len(moves)} effects={len(effects)} species={len(species)} types={len(types)}")


if __name__ == "__main__":
    sys.exit(main())
