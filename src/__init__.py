"""Import-path compatibility with the reference (`src.asr`, `src.ctc`, `src.solver`, ...): thin re-exports of the
package `end-to-end-asr-pytorch_b200` (whose directory name, fixed by the project, is not importable with a
plain `import` statement)."""
