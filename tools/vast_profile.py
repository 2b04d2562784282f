"""cProfile of one ptgs_b200.vast run (GPU box). usage: python tools/vast_profile.py [n_points] [n_cameras]"""
import cProfile, pstats, runpy, sys, io
pr = cProfile.Profile()
sys.argv = ["tools/vast_report.py"] + sys.argv[1:]
pr.enable()
try:
    runpy.run_path("tools/vast_report.py", run_name="__main__")
finally:
    pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(35)
    print(s.getvalue()[:6000])
