"""Oracle arithmetic, exact output formats, timing, tracing, post-processing."""
