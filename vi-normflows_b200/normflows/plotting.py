"""normflows/normflows/plotting.py.  Plotting is outside the hot path; matplotlib is not
installed on the build/GPU boxes, so these are import-tolerant: they draw when matplotlib
is importable and are silent no-ops otherwise."""
try:
    import matplotlib.pyplot as plt
except Exception:      # pragma: no cover - matplotlib absent
    plt = None


def plot_samples(Z, ax=None):
    if plt is None:
        return None
    plotter = plt if ax is None else ax
    dim = Z.shape[1]
    if dim == 1:
        return plotter.hist(Z, bins=25, edgecolor='k')
    elif dim == 2:
        return plotter.scatter(Z[:, 0], Z[:, 1], alpha=0.5)


def plot_obs_latent(X, Z, Xhat, Zhat):
    if plt is None:
        return None
    fig, axs = plt.subplots(ncols=2, nrows=2, sharex=True)
    for a, d, t in ((axs[0, 0], Z, 'Latent'), (axs[1, 0], Zhat, 'Variational latent'),
                    (axs[0, 1], X, 'Observed'), (axs[1, 1], Xhat, 'Variational observed')):
        plot_samples(d, a)
        a.set_title(t)


def plot_mnist(im_true, im_recon):
    if plt is None:
        return None
    fig, axs = plt.subplots(ncols=2)
    axs[0].imshow(im_true)
    axs[1].imshow(im_recon)
    axs[0].set_title('True')
    axs[1].set_title('Reconstructed')
    return axs
