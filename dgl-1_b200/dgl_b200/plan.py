"""Executor plan: the scheduler's output, in the reference's vocabulary.

The reference lowers every message-passing call to a list of executors and runs them in order
(python/dgl/runtime/ir/program.py:8-46, runtime/runtime.py:3-10).  The op names below are the
reference's (python/dgl/runtime/ir/executor.py:39-60) and `Prog.pprint()` prints the same text
as the reference's `Prog.pprint()`, so a plan made here can be diffed against the reference's
for the pairs both lower the same way (tests/test_plan.py, tests/golden/plans.json).

Each executor is a closure over torch tensors; SPMV / SPMV_WITH_DATA are the ones that launch
libdgl_b200.so kernels (autograd.aggregate).
"""

TYPE_NAMES = {"feat": "Feat", "dict": "FeatDict", "spmat": "SpMat", "idx": "Idx", "str": "Str",
              "func": "Func"}


class Var(object):
    __slots__ = ("name", "typ", "data")

    def __init__(self, name, typ, data=None):
        self.name, self.typ, self.data = name, typ, data

    def __str__(self):
        return '"%s"' % self.data if self.typ == "str" else self.name


class Exec(object):
    __slots__ = ("op", "args", "ret", "fn")

    def __init__(self, op, fn, args, ret):
        self.op, self.fn, self.args, self.ret = op, fn, args, ret

    def run(self):
        out = self.fn(*[a.data for a in self.args])
        if self.ret is not None:
            self.ret.data = out

    def __str__(self):
        argstr = ", ".join(str(a) for a in self.args)
        if self.ret is None:
            return "%s(%s)" % (self.op, argstr)
        return "%s %s = %s(%s)" % (TYPE_NAMES[self.ret.typ], self.ret.name, self.op, argstr)


class Prog(object):
    def __init__(self):
        self.execs = []
        self.varcount = 0

    def var(self, typ, data=None, name=None):
        if name is None:
            name = "_z%d" % self.varcount
            self.varcount += 1
        return Var(name, typ, data)

    def issue(self, op, fn, args, ret_typ=None):
        ret = self.var(ret_typ) if ret_typ is not None else None
        self.execs.append(Exec(op, fn, list(args), ret))
        return ret

    def lines(self):
        return [str(e) for e in self.execs]

    def pprint(self):
        print("\n".join(self.lines()))

    def run(self):
        """Runtime.run (python/dgl/runtime/runtime.py:3-10)."""
        for e in self.execs:
            e.run()
