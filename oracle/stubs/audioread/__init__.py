"""audioread stand-in (decode I/O is out of scope); exists so `import audioread.ffdec` succeeds."""
