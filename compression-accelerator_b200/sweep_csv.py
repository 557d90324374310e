"""Size-sweep results in the reference's CSV shape: `type;length;cycles;compressed-length`
(benchmark/hw_results.csv, sw_results.csv), consumable by benchmark/csv_scan.py:4-36 unchanged.

The rows come from the RoCC-shaped device object (api.RoccDevice), driven with the Scala tester's
protocol (Test.scala:17-53), i.e. the same measurement the reference harness prints per run
(`== cycles: ...; compressed length: ...`, Test.scala:53).  A "cycle" is 1 ns of device time of the
command's kernels (include/b200_rocc.h); csv_scan.py's `eff` column is then ns per input byte.

Types are spelled as csv_scan.py matches them: `random`, `real`, `repeat` (csv_scan.py:22-27)."""
import csv
import io

SWEEP_SIZES = (10, 20, 50, 100, 200, 500, 1000, 2000, 5000, 10000, 20000, 50000)   # DataGenerator.scala:27
TYPES = ("random", "real", "repeat")
HEADER = "type;length;cycles;compressed-length"


def sweep_rows(run, data_of, sizes=SWEEP_SIZES, types=TYPES, repeats=3):
    """run(data: bytes) -> (cycles, compressed_length); data_of(type, n) -> bytes.
    Best-of-`repeats` cycles per point (the first command also pays one-time set-up)."""
    rows = []
    for typ in types:
        for n in sizes:
            data = data_of(typ, n)
            best, clen = None, None
            for _ in range(max(1, repeats)):
                cycles, got = run(data)
                if clen is not None and got != clen:
                    raise RuntimeError("compressed length changed between runs: %s %d" % (typ, n))
                clen = got
                best = cycles if best is None else min(best, cycles)
            rows.append((typ, n, int(best), int(clen)))
    return rows


def to_csv(rows) -> str:
    out = io.StringIO()
    out.write(HEADER + "\n")
    for typ, n, cycles, clen in rows:
        out.write("%s;%d;%d;%d\n" % (typ, n, cycles, clen))
    return out.getvalue()


def scan(text: str):
    """What benchmark/csv_scan.py:9-27 derives per row: {type: {length: (ratio, cycles_per_byte)}}."""
    res = {}
    for rowf in csv.reader(io.StringIO(text), delimiter=";"):
        row = [c for c in rowf if c]
        if len(row) < 2 or row[0] == "type":
            continue
        typ, length, cycles, clen = row[0], int(row[1]), float(row[2]), int(row[3])   # csv_scan.py:16-20 uses float()
        res.setdefault(typ, {})[length] = (length / clen, cycles / length)
    return res


def device_runner(dev):
    """run() over an api.RoccDevice: load at address 0, SET_LENGTH + COMPRESS, read cycles and length."""
    def run(data):
        dev.load_mem(data, 0)
        cycles, clen, _dst = dev.run_compress(len(data))
        return cycles, clen
    return run
