"""Generates tests/golden/gp_fit/*: what the REFERENCE's GP_fit.create_GP writes and returns when its numerics
(len_scale_opt, fill_convex_hull, GP) and its input (read_yaml) are replaced by the deterministic stand-ins of
tests/gp_fit_stubs.py.  The orchestration under test (experiment loop, column naming, the three output modes, the outer
merges with +inf fill, CSV text) is the reference's own code, imported from /root/reference.  Run in the build container:

    python tests/golden/make_gp_fit_golden.py
"""
import contextlib, io, json, os, shutil, sys, tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import gp_fit_stubs as stubs

OUT = os.path.join(HERE, "gp_fit")


def main():
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import ref_shim
    ref_shim.install_smt_stub()              # find_len_scales imports smt at module top; len_scale_opt is replaced anyway
    sys.path.insert(0, "/root/reference")
    import GP_fit as ref
    os.makedirs(OUT, exist_ok=True)
    manifest = {}
    for mode in stubs.MODES:
        work = tempfile.mkdtemp()
        cwd = os.getcwd()
        os.chdir(work)
        try:
            stubs.install(ref, mode)
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                df, nd = ref.create_GP()
            files = {}
            for root, _, names in os.walk(work):
                for n in names:
                    rel = os.path.relpath(os.path.join(root, n), work)
                    files[rel] = open(os.path.join(root, n)).read()
            manifest[mode] = {"stdout": buf.getvalue(), "ndims": nd, "returned_csv": df.to_csv(index=False), "files": files}
        finally:
            os.chdir(cwd)
            shutil.rmtree(work)
    with open(os.path.join(OUT, "create_gp_cases.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    print("wrote", os.path.join(OUT, "create_gp_cases.json"), {m: len(v["files"]) for m, v in manifest.items()})


if __name__ == "__main__":
    main()
