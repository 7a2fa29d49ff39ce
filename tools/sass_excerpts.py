"""Committed SASS evidence per fused kernel: for each kernel of libtfdl_kernels.so named below, the instruction
neighbourhoods around its Blackwell-specific instructions (UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG /
UTMAREDG = TMA tensor load / store / reduce, SYNCS = mbarrier, MULTIMEM / .MC = NVLS, .SYS loads and stores on peer
pointers), written to profiles/sass/<kernel>.sass. The census is profiles/sass_summary.md (tools/sass_summary.py).

    python tools/sass_excerpts.py            # after tools/build_native.py"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tensorflow-distributed-learning_b200", "_native", "libtfdl_kernels.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KERNELS = ["conv_igemm_kernel<128, 6>", "tma_gemm_pair_kernel<256, 6>", "conv12_fwd_tma_kernel<13, 5>", "conv21_bwd_tma_kernel<13>",
           "conv2_wgrad_tma_kernel<13, 4>", "dense_block_kernel", "attention_fwd_kernel", "fused_ar_push_kernel",
           "fused_ar_push2_kernel", "fused_ar_oneshot_kernel", "fused_ar_sliced_kernel<true, 4>",
           "fused_ar_sliced_kernel<false, 2>", "allreduce_sum_kernel<true>", "stage_inputs_kernel"]
MARK = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|UTMALDG|UTMASTG|UTMAREDG|UTMACMDFLUSH|UTCBAR|SYNCS|MULTIMEM|LDG\S*\.SYS|STG\S*\.SYS|"
                  r"LD\.E\S*\.SYS|ST\.E\S*\.SYS|REDG?\S*|ATOMG?\S*|ACQBULK|UBLKCP)")
CONTEXT, MAX_LINES = 6, 320


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
    funcs, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            funcs[cur].append(line.rstrip())
    names = subprocess.run(["c++filt"], input="\n".join(funcs), stdout=subprocess.PIPE, text=True).stdout.splitlines()
    os.makedirs(OUT, exist_ok=True)
    for mangled, name in zip(funcs, names):
        short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "")).replace("void ", "").replace("tfdl::", "")
        if short not in KERNELS:
            continue
        body = [l for l in funcs[mangled] if re.search(r"/\*[0-9a-f]{4}\*/", l)]  # instruction lines only
        hits = [i for i, l in enumerate(body) if MARK.search(l)]
        keep = set()
        for i in hits:
            keep.update(range(max(0, i - CONTEXT), min(len(body), i + CONTEXT + 1)))
        lines, last = [], -2
        for i in sorted(keep):
            if i != last + 1:
                lines.append("        ...")
            lines.append(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", body[i]))
            last = i
            if len(lines) >= MAX_LINES:
                lines.append("        ... (truncated)")
                break
        counts = {}
        for i in hits:
            k = MARK.search(body[i]).group(1).split(".")[0]
            counts[k] = counts.get(k, 0) + 1
        fn = os.path.join(OUT, re.sub(r"[^A-Za-z0-9_]+", "_", short).strip("_") + ".sass")
        with open(fn, "w") as f:
            f.write(f"// {name}\n// {len(body)} SASS instructions; marked: {counts}\n// cuobjdump -sass of libtfdl_kernels.so (sm_100a), "
                    f"+-{CONTEXT} instructions around each marked one\n")
            f.write("\n".join(lines) + "\n")
        print(f"{short}: {len(body)} instructions, {len(hits)} marked -> {os.path.relpath(fn, ROOT)}")


if __name__ == "__main__":
    main()
