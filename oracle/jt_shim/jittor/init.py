"""jittor.init stand-in: nothing on the post-processing path needs it."""
