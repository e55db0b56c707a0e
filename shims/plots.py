"""`plots.plot_token_probability` (src/plots.py:4): heat map of one token's probability over layers x
positions from a dense ``all_probs`` [L, T, V] array. matplotlib is imported when the function is called,
so the drivers import without it."""


def plot_token_probability(all_probs, token_id, tokenizer, input_words, figsize=(22, 11), start_idx=0,
                           font_size=30, title_font_size=36, tick_font_size=32, colormap="viridis"):
    import matplotlib.pyplot as plt

    grid = all_probs[:, start_idx:, token_id]            # [layers, positions]
    words = list(input_words[start_idx:])
    fig, ax = plt.subplots(figsize=figsize)
    plt.rcParams.update({"font.size": font_size})
    image = ax.imshow(grid, cmap=colormap, aspect="auto", vmin=0, vmax=1, interpolation="nearest")
    bar = fig.colorbar(image, ax=ax)
    bar.ax.tick_params(labelsize=tick_font_size)
    ax.set_ylabel("Layers", fontsize=title_font_size)
    ax.set_yticks(list(range(grid.shape[0]))[::4])
    ax.tick_params(axis="y", labelsize=tick_font_size)
    if words:
        ax.set_xticks(list(range(len(words))))
        ax.set_xticklabels(words, rotation=75, ha="right", fontsize=font_size)
    plt.tight_layout()
    return fig
