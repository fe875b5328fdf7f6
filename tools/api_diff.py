"""Public API of the reference (installed under baseline/_ref) vs nanofed_b200: every public name of
every ``nanofed.*`` module must exist in the matching ``nanofed_b200.*`` module, and every method /
function must accept the reference's parameter names.  Exit code 1 on any gap."""
import importlib, inspect, pkgutil, sys
ROOT = __import__("pathlib").Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
sys.path.insert(0, str(ROOT))
import nanofed
mods = [m.name for m in pkgutil.walk_packages(nanofed.__path__, "nanofed.")]
missing, sigdiff = [], []
for name in ["nanofed"] + mods:
    try:
        ref = importlib.import_module(name)
    except Exception as e:
        print("ref import fail", name, e); continue
    mine_name = name.replace("nanofed", "nanofed_b200", 1)
    try:
        mine = importlib.import_module(mine_name)
    except Exception as e:
        missing.append((name, "MODULE", str(e)[:60])); continue
    pub = getattr(ref, "__all__", None) or [n for n in dir(ref) if not n.startswith("_")]
    for n in pub:
        r = getattr(ref, n, None)
        if inspect.ismodule(r): continue
        if getattr(r, "__module__", "").split(".")[0] not in ("nanofed",): continue
        m = getattr(mine, n, None)
        if m is None:
            missing.append((name, n, "")); continue
        if inspect.isclass(r):
            for an, av in inspect.getmembers(r):
                if an.startswith("_") and an != "__init__": continue
                if not (inspect.isfunction(av) or isinstance(av, property)): continue
                mv = getattr(m, an, None)
                if mv is None:
                    missing.append((name, f"{n}.{an}", "")); continue
                if inspect.isfunction(av) and inspect.isfunction(mv):
                    try:
                        rs, ms = inspect.signature(av), inspect.signature(mv)
                    except Exception: continue
                    rp = [p for p in rs.parameters.values()]
                    mp = {p.name: p for p in ms.parameters.values()}
                    for p in rp:
                        if p.name not in mp and p.kind not in (p.VAR_KEYWORD, p.VAR_POSITIONAL):
                            sigdiff.append((name, f"{n}.{an}", f"param {p.name} missing: ref{rs} mine{ms}"))
        elif inspect.isfunction(r) and inspect.isfunction(m):
            rs, ms = inspect.signature(r), inspect.signature(m)
            for p in rs.parameters.values():
                if p.name not in ms.parameters and p.kind not in (p.VAR_KEYWORD, p.VAR_POSITIONAL):
                    sigdiff.append((name, n, f"param {p.name} missing: ref{rs} mine{ms}"))
print("MISSING:"); [print("  ", x) for x in missing]
print("SIGDIFF:"); [print("  ", x) for x in sigdiff]
sys.exit(1 if (missing or sigdiff) else 0)
