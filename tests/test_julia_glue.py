"""julia/KirstineB200.jl cannot be executed here (no Julia toolchain in the container or on the GPU box), so its
binding to the C ABI is checked mechanically: every `ccall` must name a function include/kirstine_b200.h declares,
pass as many values as it declares types, declare as many types as the prototype has parameters, and use a Julia
type compatible with each C type; the Julia mirror of kst_exchange_options must match the C struct field by field;
and the glue must cover the entry points the reference's public API for this path needs.  A drifting signature on
either side fails here."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kirstine_b200.h")
GLUE = os.path.join(ROOT, "julia", "KirstineB200.jl")

# C parameter type (normalised) -> Julia types a ccall may declare for it
COMPAT = {
    "int": {"Cint"},
    "int64_t": {"Int64", "Clonglong"},
    "uint64_t": {"UInt64"},
    "double": {"Cdouble", "Float64"},
    "double*": {"Ptr{Cdouble}", "Ref{Cdouble}", "Ref{Float64}", "Ptr{Float64}"},
    "int*": {"Ptr{Cint}", "Ref{Cint}"},
    "int64_t*": {"Ptr{Int64}", "Ref{Int64}"},
    "void*": {"Ptr{Cvoid}"},
    "kst_problem*": {"Ptr{Cvoid}"},
    "kst_problem**": {"Ref{Ptr{Cvoid}}", "Ptr{Ptr{Cvoid}}"},
    "kst_exchange_options*": {"Ref{ExchangeOptions}", "Ptr{ExchangeOptions}"},
    "char*": {"Cstring", "Ptr{UInt8}"},
}
FIELD = {"int": "Cint", "double": "Cdouble", "uint64_t": "UInt64", "int64_t": "Int64"}


def _strip_comments(text):
    return re.sub(r"/\*.*?\*/", " ", text, flags=re.S)


def _ctype(decl, with_name=True):
    """'const double *points' -> 'double*'."""
    d = decl.replace("const", " ").strip()
    stars = d.count("*")
    d = d.replace("*", " ")
    words = d.split()
    if with_name and len(words) > 1:
        words = words[:-1]
    return " ".join(words) + "*" * stars


def header_prototypes():
    text = " ".join(_strip_comments(open(HEADER).read()).split())
    protos = {}
    for m in re.finditer(r"KST_API ([\w ]+?\*?) ?(\**)(kst_\w+)\(([^)]*)\);", text):
        ret = _ctype(m.group(1) + m.group(2), with_name=False)
        args = m.group(4).strip()
        params = [] if args in ("", "void") else [_ctype(a) for a in args.split(",")]
        protos[m.group(3)] = (ret, params)
    return protos


def _split_top(s):
    """Split at top-level commas (parentheses, brackets and braces nest)."""
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def glue_ccalls():
    text = re.sub(r"#.*", "", open(GLUE).read())
    calls = []
    for m in re.finditer(r"ccall\(", text):
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(text[i], 0)
            i += 1
        parts = _split_top(text[m.end():i - 1])
        name = re.match(r"\(:(\w+), LIB\)", parts[0]).group(1)
        types = _split_top(parts[2].strip()[1:-1])
        calls.append((name, parts[1], types, parts[3:]))
    return calls


def test_every_ccall_matches_the_header():
    protos = header_prototypes()
    assert len(protos) >= 30 and "kst_exchange_run" in protos and "kst_pso_optimize" in protos
    calls = glue_ccalls()
    assert len(calls) >= 10
    for name, ret, types, values in calls:
        assert name in protos, "ccall to %s, which the header does not declare" % name
        cret, cparams = protos[name]
        assert ret in COMPAT[cret], "%s: return type %s vs C %s" % (name, ret, cret)
        assert len(types) == len(cparams), "%s: %d Julia argument types, %d C parameters" % (name, len(types), len(cparams))
        assert len(values) == len(types), "%s: %d values for %d declared types" % (name, len(values), len(types))
        for k, (jt, ct) in enumerate(zip(types, cparams)):
            assert jt in COMPAT[ct], "%s, argument %d: Julia %s vs C %s" % (name, k + 1, jt, ct)


def test_exchange_options_struct_mirrors_the_header():
    h = _strip_comments(open(HEADER).read())
    body = re.search(r"typedef struct kst_exchange_options \{(.*?)\} kst_exchange_options;", h, flags=re.S).group(1)
    cfields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if decl:
            ty, names = decl.split(None, 1)
            cfields += [(n.strip(), FIELD[ty]) for n in names.split(",")]
    j = re.sub(r"#.*", "", open(GLUE).read())
    jbody = re.search(r"struct ExchangeOptions\n(.*?)\nend", j, flags=re.S).group(1)
    jfields = [tuple(x.strip() for x in line.split("::")) for line in jbody.strip().splitlines()]
    assert jfields == cfields


def test_glue_covers_the_reference_api_of_the_path():
    used = {c[0] for c in glue_ccalls()}
    # objective / efficiency / shannon_information, gateauxderivative, informationmatrix, solve with
    # DirectMaximization and with Exchange, RegularizedFisherMatrix, error strings, handle life cycle
    for need in ("kst_problem_create", "kst_problem_destroy", "kst_last_error", "kst_objective_batch", "kst_gateaux_prepare",
                 "kst_gateaux_eval", "kst_informationmatrix", "kst_pso_optimize", "kst_exchange_run",
                 "kst_problem_set_regularization"):
        assert need in used, need
    j = open(GLUE).read()
    for ident in ("DTRMod", "CopyBoth", "FixedDose", "FixedTimes", "EquidistantTimes", "LogEquidistantTimes", "VUModAR1",
                  "solve_with(dp::DesignProblem{<:Any,<:Any,<:Registered}, strategy::Exchange{Pso,Pso}"):
        assert ident in j, ident
    # the model / criterion / covariate-parameterization ids in the glue are the header's
    h = open(HEADER).read()
    ids = dict(re.findall(r"#define (KST_(?:MODEL|CP|CRIT)_\w+) (\d+)", h))
    for jl, macro in (("model_id(::EmaxModel) = 1", "KST_MODEL_EMAX"), ("model_id(::SigEmaxModel) = 2", "KST_MODEL_SIGEMAX"),
                      ("model_id(::TPCModel) = 3", "KST_MODEL_TPC"), ("model_id(::VUMod) = 4", "KST_MODEL_VU"),
                      ("model_id(::DTRMod) = 5", "KST_MODEL_DTR"), ("model_id(::LinLogReg) = 6", "KST_MODEL_LINLOGREG"),
                      ("model_id(::SinLogReg) = 7", "KST_MODEL_SINLOGREG"), ("model_id(::VUModAR1) = 8", "KST_MODEL_VUAR"),
                      ("crit_id(::DexpCriterion) = 2", "KST_CRIT_DEXP"), ("cp_spec(::DTRMod, ::CopyBoth) = (2,", "KST_CP_DTR_COPYBOTH"),
                      ("cp_spec(::DTRMod, cp::LogEquidistantTimes) = (6,", "KST_CP_DTR_LOGEQUIDISTANT")):
        assert jl in j and jl.rstrip(",").rsplit("(", 1)[-1].rsplit(" ", 1)[-1].strip("(,") == ids[macro], (jl, macro)
