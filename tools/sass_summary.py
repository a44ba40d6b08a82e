import collections, re, subprocess
out = subprocess.run(["cuobjdump", "-sass", "graphnet_b200/csrc/libgraphnet_b200.so"], capture_output=True, text=True).stdout
cur = None
stats = collections.defaultdict(lambda: collections.Counter())
inst = collections.Counter()
pats = {"UTCHMMA": r"\bUTCHMMA", "2CTA": r"UTCHMMA\.2CTA", "UTCBAR": r"\bUTCBAR", "LDTM": r"\bLDTM", "STTM": r"\bSTTM",
        "UTMALDG": r"\bUTMALDG", "MULTICAST": r"UTMALDG\S*MULTICAST", "UTMASTG": r"\bUTMASTG", "SYNCS": r"\bSYNCS", "FADD2": r"\bFADD2",
        "HFMA2.RELU": r"HFMA2\.RELU", "RED": r"\b(REDG|ATOMG|RED\.|ATOM\.)"}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"<.*", "", re.sub(r"^void ", "", name)).split("(")[0]
        inst[cur] += 1
        continue
    if cur and re.search(r"/\*[0-9a-f]{4}\*/", line):
        stats[cur]["instr"] += 1
        for k, p in pats.items():
            if re.search(p, line):
                stats[cur][k] += 1
print("# Round 2, end state - SASS instruction counts of the built library (`cuobjdump -sass graphnet_b200/csrc/libgraphnet_b200.so`, sm_100a)\n")
print("Static counts per kernel (instances of a template are summed).  `UTCHMMA` = tcgen05.mma, `.2CTA` its cta_group::2 form, `UTCBAR` = tcgen05.commit, `LDTM` / `STTM` = tcgen05.ld / st, `UTMALDG` = TMA tensor load (`.MULTICAST`: cluster multicast), `UTMASTG` = TMA tensor store, `SYNCS` = mbarrier, `FADD2` = packed fp32 add.\n")
cols = ["UTCHMMA", "2CTA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "MULTICAST", "UTMASTG", "SYNCS", "FADD2", "HFMA2.RELU", "RED"]
print("| kernel | instances | SASS instr. | " + " | ".join(cols) + " |")
print("|---|---:|---:|" + "---:|" * len(cols))
for k, c in sorted(stats.items(), key=lambda kv: -kv[1]["instr"]):
    print(f"| `{k}` | {inst[k]} | {c['instr']} | " + " | ".join(str(c[x]) for x in cols) + " |")
