def nvtx_scope(name):
    def deco(fn):
        return fn
    return deco
