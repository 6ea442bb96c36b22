"""bucketmap_analyze: the metrics of the reference's benchmark/sam_file_analyzer.cpp:250-358.

The reference's tool has its input paths compiled in (main(), :371-383) and reads SAM through SeqAn3, so it cannot be run
here; the expected numbers below are worked out by hand from its rules, and a second, independent restatement in Python
checks them on a random case."""
import json
import os
import random
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "bucket-map_b200", "bin", "bucketmap_analyze")


@pytest.fixture(scope="module", autouse=True)
def _built():
    p = subprocess.run(["make", "-s", "-C", ROOT, "bucket-map_b200/bin/bucketmap_analyze"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr


def _run(*args):
    p = subprocess.run([EXE, *args], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    return p.stdout


HEADER = "@HD\tVN:1.6\n@SQ\tSN:chrA\tLN:1000\n@SQ\tSN:chrB\tLN:1000\n"


def _rec(name, flag, ref, pos):
    return "%s\t%d\t%s\t%d\t60\t*\t*\t0\t0\tACGT\tEEEE\n" % (name, flag, ref, pos)


def test_hand_worked_case(tmp_path):
    fq = tmp_path / "r.fastq"
    fq.write_text("".join("@%s\nACGT\n+\nEEEE\n" % n for n in ["r0", "r1 extra", "r2/1", "r3", "r4"]))
    truth = tmp_path / "t.position_ground_truth"
    #            origin pos rev
    truth.write_text("0 101 0 4=\n1 501 1 4=\n0 301 0 4=\n1 11 0 4=\n0 900 1 4=\n")
    sam = tmp_path / "o.sam"
    sam.write_text(HEADER +
                   _rec("r0", 0, "chrA", 101) +        # pos0 = 100 vs 101: within tolerance -> correct
                   _rec("r1_extra", 16, "chrB", 520) +  # pos0 = 519 vs 501: off by 18 -> wrong (tolerance 10)
                   _rec("r2", 0, "chrA", 311) +        # 310 vs 301 -> correct; name cut at '/'
                   _rec("r2", 0, "chrB", 311) +        # second record of r2: not unique -> r2 leaves the read counts
                   _rec("r3", 4, "*", 0) +             # unmapped
                   _rec("zz", 0, "chrA", 5) +          # unknown read: ignored
                   _rec("r4", 0, "chrA", 900))         # wrong strand
    m = json.loads(_run("--reads", str(fq), "--truth", str(truth), "--json", str(sam)))[0]
    assert m["reads"] == 5 and m["random_reads"] == 0 and m["answer_unique"] == 5
    assert m["mapped_locations"] == 5 and m["acceptable_locations"] == 2
    assert m["unique"] == [4, 1, 0, 0]
    assert m["mapped_reads"] == 3          # r0, r1, r4 (r2 is not uniquely mapped by the tested file, r3 unmapped)
    assert m["correct_reads"] == 1         # r0 (r2's correct record does not count: :306-314)
    text = _run("--reads", str(fq), "--truth", str(truth), str(sam))
    assert "[BENCHMARK]\tTotal number of reads: 5.\n" in text
    assert "4 1 0 0.\n" in text
    assert "[BENCHMARK]\tTotal number of mapped reads: 3 (60%).\n" in text
    assert "[BENCHMARK]\tTotal number of correctly mapped reads (sensitivity): 1 (20%).\n" in text
    assert "[BENCHMARK]\tNumber of mapped locations returned: 5 (1.66667 per mapped read).\n" in text
    assert "[BENCHMARK]\tTotal number of acceptable mapped locations (precision): 2 (precision: 40%).\n" in text
    # tolerance is an argument (the class default is 5, main() uses 10)
    m20 = json.loads(_run("--reads", str(fq), "--truth", str(truth), "--tolerance", "20", "--json", str(sam)))[0]
    assert m20["acceptable_locations"] == 3 and m20["correct_reads"] == 2


def test_best_alignment_file_as_answer_and_directory(tmp_path):
    fq = tmp_path / "r.fastq"
    fq.write_text("".join("@q%d\nACGT\n+\nEEEE\n" % i for i in range(4)))
    best = tmp_path / "best.sam"
    best.write_text(HEADER + _rec("q0", 0, "chrA", 10) + _rec("q1", 16, "chrB", 20) + _rec("q1", 0, "chrA", 700) +
                    _rec("q2", 4, "*", 0))
    d = tmp_path / "out"
    d.mkdir()
    (d / "a.sam").write_text(HEADER + _rec("q0", 0, "chrA", 12) + _rec("q1", 0, "chrA", 705) + _rec("q3", 0, "chrA", 1))
    (d / "b.sam").write_text(HEADER)
    (d / "note.txt").write_text("not a sam file")
    out = json.loads(_run("--reads", str(fq), "--best", str(best), "--json", str(d)))
    assert [os.path.basename(m["sam"]) for m in out] == ["a.sam", "b.sam"]
    a, b = out
    assert a["answer_unique"] == 1                      # q0 only: q1 has two answers, q2 / q3 none
    assert a["unique"] == [1, 0, 3, 0]
    assert a["mapped_reads"] == 3 and a["correct_reads"] == 2 and a["acceptable_locations"] == 2
    assert b["mapped_reads"] == 0 and b["mapped_locations"] == 0


def test_dwgsim_names(tmp_path):
    fa = tmp_path / "g.fa"
    fa.write_text(">NC_1 some text\nACGT\n>NC_2\nACGT\n")
    fq = tmp_path / "r.fastq"
    # contig_start_start2_strand1_strand2_random1_random2_...:...  -> parts[2] start, parts[4] strand, parts[6] random
    names = ["NC_2_150_300_1_0_0_0_1:0:0_0:0:0_0/1", "NC_1_40_200_0_1_0_0_0:0:0_0:0:0_1/1", "rand_0_7_9_0_0_1_1_0:0:0_0:0:0_2/1"]
    fq.write_text("".join("@%s\nACGT\n+\nEEEE\n" % n for n in names))
    sam = tmp_path / "o.sam"
    sam.write_text("@SQ\tSN:NC_1\tLN:9\n@SQ\tSN:NC_2\tLN:9\n" + _rec(names[0][:-2], 16, "NC_2", 155) +
                   _rec(names[1][:-2], 0, "NC_1", 60) + _rec(names[2][:-2], 0, "NC_1", 1))
    m = json.loads(_run("--reads", str(fq), "--dwgsim", "--genome", str(fa), "--json", str(sam)))[0]
    assert m["reads"] == 3 and m["random_reads"] == 1 and m["mapped_random"] == 1
    assert m["mapped_reads"] == 3 and m["correct_reads"] == 1 and m["acceptable_locations"] == 1


def _restate(n, answers, records, tol):
    mapped, uniq, correct = [False] * n, [True] * n, [False] * n
    locs = acc = 0
    for r, rev, ref, pos in records:
        if mapped[r]:
            uniq[r] = False
        mapped[r] = True
        locs += 1
        if any(rev == a[0] and ref == a[1] and abs(pos - a[2]) <= tol for a in answers[r]):
            correct[r] = True
            acc += 1
    return (sum(m and u for m, u in zip(mapped, uniq)), sum(c and u for c, u in zip(correct, uniq)), locs, acc)


def test_random_case_against_python_restatement(tmp_path):
    rng = random.Random(5)
    n = 400
    fq = tmp_path / "r.fastq"
    fq.write_text("".join("@%d\nACGT\n+\nEEEE\n" % i for i in range(n)))
    truth = [(rng.random() < 0.5, rng.randrange(2), rng.randrange(1, 900)) for _ in range(n)]
    (tmp_path / "t.ground_truth").write_text("".join("%d %d %d 4=\n" % (ref, pos, rev) for rev, ref, pos in truth))
    records, lines = [], []
    for i in range(n):
        for _ in range(rng.choice([0, 1, 1, 1, 2, 3])):
            rev, ref, pos = truth[i]
            if rng.random() < 0.3:
                rev = not rev
            if rng.random() < 0.2:
                ref = 1 - ref
            pos0 = max(0, pos + rng.randrange(-15, 16))
            records.append((i, rev, ref, pos0))
            lines.append(_rec(str(i), 16 if rev else 0, "chrA" if ref == 0 else "chrB", pos0 + 1))
    (tmp_path / "o.sam").write_text(HEADER + "".join(lines))
    m = json.loads(_run("--reads", str(fq), "--truth", str(tmp_path / "t.ground_truth"), "--json", str(tmp_path / "o.sam")))[0]
    want = _restate(n, [[t] for t in truth], records, 10)
    assert (m["mapped_reads"], m["correct_reads"], m["mapped_locations"], m["acceptable_locations"]) == want


def test_missing_arguments_fail():
    p = subprocess.run([EXE, "--reads", "x.fastq"], capture_output=True, text=True)
    assert p.returncode == 2 and "[ERROR]" in p.stderr
