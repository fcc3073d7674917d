from .run_pipeline import main

main()
