"""gin-config is optional: the reference decorates its solvers with `@gin.configurable()`
(`fmap2pointmap_solvers/naive.py:5`, `spacial_filtering.py:4`, `__init__.py:6`) so that
`config/dpfm_orig.gin:71` can bind `choose_fmap2pointmap_solver.solver = @spacial_filtering_fmap2pointmap`.
When gin is importable the same names are registered; otherwise the decorator is a no-op."""
try:
    import gin  # type: ignore

    def configurable(fn):
        try:
            return gin.configurable()(fn)
        except Exception:  # already registered (module reloaded) or a stub gin
            return fn
except Exception:  # gin not installed
    gin = None

    def configurable(fn):
        return fn
