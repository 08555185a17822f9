# placeholder, filled below
