import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pinyon_b200 import Context
ctx = Context(0)
names = ["LOP3", "LOP3+SHF+IADD3", "LOP3+SHF+IMAD", "IMAD", "IMAD.HI", "LOP3:IMAD 1:1", "LOP3:IMAD.HI(:IMAD)",
         "IMAD.WIDE", "LOP3:IMAD.WIDE 2:1", "SHF"]
for v, nm in enumerate(names):
    r = ctx.int_peak(v, 5)
    print(f"variant {v} {nm:22s} {r/1e12:7.2f} Top/s  = {r/(148*1.965e9):6.1f} lanes/clk/SM")
