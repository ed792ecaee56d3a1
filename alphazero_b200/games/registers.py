"""Name -> class maps with the reference's keys (alphazero/games/registers.py:11-50); Othello and Connect4 are the
games the B200 engine implements (TicTacToe stays on the reference's Python path)."""
from ..base import DataTransf, MoveFormat, dotdict
from .connect4 import Connect4Board, Connect4Config, Connect4Net
from .othello import OthelloBoard, OthelloConfig, OthelloNet

GAMES_SET = {"othello", "connect4"}
CONFIGS_REGISTER = {"othello": OthelloConfig, "connect4": Connect4Config}
BOARDS_REGISTER = {"othello": OthelloBoard, "connect4": Connect4Board}
NETWORKS_REGISTER = {"othello": OthelloNet, "connect4": Connect4Net}
MOVE_FORMATS_REGISTER = {"othello": MoveFormat.ROW_COL, "connect4": MoveFormat.COL}
DATA_AUGMENT_STRATEGIES = {
    "othello": dotdict({"reflection": DataTransf.REFLECT_H,
                        "rotations": [DataTransf.ROTATE_90, DataTransf.ROTATE_180, DataTransf.ROTATE_270]}),
    "connect4": dotdict({"reflection": DataTransf.REFLECT_H, "rotations": []}),
}
