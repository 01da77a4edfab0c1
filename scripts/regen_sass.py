"""Regenerate the committed SASS listings (profiles/sass/) from the objects of the current build and fail when a
listing is older than the source it documents.

    python scripts/regen_sass.py            write one <source>.sass per csrc/*.cu + MNEMONICS.txt
    python scripts/regen_sass.py --check    exit 1 if a listing is missing or its recorded source hash is stale

Every listing starts with '// source: csrc/<file>.cu sha256=<hash>' so staleness is a property of the file itself,
not of timestamps.  MNEMONICS.txt counts the Blackwell-specific instructions per source (tensor core UTCHMMA / TMEM
LDTM, TMA UTMALDG / UBLKCP / UBLKRED, multimem LDGMC / STGMC, vector REDG, mbarrier SYNCS) — the evidence the profiling
recipe asks for."""
import collections
import hashlib
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
OUT = ROOT / "profiles" / "sass"
INTERESTING = re.compile(r"^(UTC\w+|UTMA\w+|UBLK\w+|LDTM|STTM|LDGMC|STGMC|REDGMC|REDG|ATOMG?|SYNCS|ELECT|LDGSTS|UTCBAR|MULTIMEM|FENCE|MEMBAR)")


def digest(src: Path) -> str:
    h = hashlib.sha256(src.read_bytes())
    for hdr in sorted(list((ROOT / "csrc").glob("*.cuh")) + list((ROOT / "csrc").glob("*.h")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()[:16]


def main() -> int:
    check = "--check" in sys.argv
    srcs = sorted((ROOT / "csrc").glob("*.cu"))
    if check:
        bad = []
        for src in srcs:
            f = OUT / (src.stem + ".sass")
            first = f.read_text().splitlines()[0] if f.exists() else ""
            if f"sha256={digest(src)}" not in first:
                bad.append(src.name)
        if bad:
            print("stale or missing SASS listings for:", ", ".join(bad), "— run python scripts/regen_sass.py")
            return 1
        print("SASS listings are current")
        return 0
    from distributed_tensorflow_models_b200._native import build as nb
    nb.build()
    OUT.mkdir(parents=True, exist_ok=True)
    lines = ["Blackwell-specific SASS mnemonics per source (cuobjdump -sass of the sm_100a objects; scripts/regen_sass.py)"]
    for src in srcs:
        obj = nb.OBJ_DIR / (src.stem + ".o")
        sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True, check=True).stdout
        # keep the listing readable: drop the encoding comments
        body = re.sub(r"\s*/\* 0x[0-9a-f]+ \*/", "", sass)
        (OUT / (src.stem + ".sass")).write_text(f"// source: csrc/{src.name} sha256={digest(src)}\n" + body)
        counts = collections.Counter()
        n = 0
        for ln in body.splitlines():
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", ln)
            if not m:
                continue
            n += 1
            op = m.group(1)
            if INTERESTING.match(op):
                counts[".".join(op.split(".")[:4])] += 1
        kernels = len(re.findall(r"^\s*Function : ", body, flags=re.M))
        lines.append(f"{src.stem}.sass: {kernels} kernels, {n} instructions; " +
                     ", ".join(f"{k} x{v}" for k, v in sorted(counts.items())))
    (OUT / "MNEMONICS.txt").write_text("\n".join(lines) + "\n")
    print("wrote", len(srcs), "listings to", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
