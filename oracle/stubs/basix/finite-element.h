// Empty stand-in: the reference CPU kernels include this Basix header but use nothing from it.
