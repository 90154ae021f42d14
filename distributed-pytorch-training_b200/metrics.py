"""Logging sinks with the reference's exact stdout / CSV formats (train_ddp.py:229-244,
326-327, 349-355, 374-384) plus an optional JSON sink for the device-timed fields."""
from __future__ import annotations

import json
from pathlib import Path

CSV_HEADER = "epoch,train_loss,train_acc,val_loss,val_acc,epoch_time_seconds\n"


def banner(device, world_size: int, amp: bool) -> str:
    return f"Using device: {device}, world_size={world_size}, amp={amp}"


def step_line(epoch: int, step: int, steps_per_epoch: int, avg_loss: float, avg_acc: float,
              throughput: float) -> str:
    return (f"Epoch [{epoch + 1}] Step [{step}/{steps_per_epoch}] "
            f"Loss: {avg_loss:.4f}  "
            f"Acc: {avg_acc:.2f}%  "
            f"Throughput: {throughput:.2f} samples/s (global)")


def epoch_line(epoch: int, epochs: int, train_loss: float, train_acc: float, val_loss: float,
               val_acc: float, epoch_time: float) -> str:
    return (f"[Epoch {epoch + 1}/{epochs}] "
            f"Train: loss={train_loss:.4f}, acc={train_acc:.2f}% | "
            f"Val: loss={val_loss:.4f}, acc={val_acc:.2f}% | "
            f"Epoch time: {epoch_time:.2f}s")


def csv_row(epoch: int, train_loss: float, train_acc: float, val_loss: float, val_acc: float,
            epoch_time: float) -> str:
    return (f"{epoch + 1},{train_loss:.4f},{train_acc:.2f},"
            f"{val_loss:.4f},{val_acc:.2f},{epoch_time:.4f}\n")


class MetricsCSV:
    """``metrics_rank0.csv``: header only when the file is absent, rows appended across runs
    (reference quirk kept on purpose, train_ddp.py:350-355)."""

    def __init__(self, output_dir: str, enabled: bool = True) -> None:
        self.path = Path(output_dir) / "metrics_rank0.csv"
        self.enabled = enabled
        if enabled and not self.path.exists():
            self.path.parent.mkdir(parents=True, exist_ok=True)
            with self.path.open("w") as f:
                f.write(CSV_HEADER)

    def append(self, epoch, train_loss, train_acc, val_loss, val_acc, epoch_time) -> None:
        if not self.enabled:
            return
        with self.path.open("a") as f:
            f.write(csv_row(epoch, train_loss, train_acc, val_loss, val_acc, epoch_time))


class MetricsJSON:
    def __init__(self, path: str, enabled: bool = True) -> None:
        self.path = Path(path) if path else None
        self.enabled = enabled and bool(path)
        self.rows: list[dict] = []

    def append(self, row: dict) -> None:
        if not self.enabled:
            return
        self.rows.append(row)
        self.path.parent.mkdir(parents=True, exist_ok=True)
        with self.path.open("w") as f:
            json.dump(self.rows, f, indent=1)
